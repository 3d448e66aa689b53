// Micro-benchmark: does the value slab's layout limit the backward kernel?  Each CTA of 128 threads owns 128 paths and walks
// the G inventory rows in order, reading one 1 KB row segment of `in` and writing one of `out` per step (the backward
// kernel's DRAM traffic with all arithmetic removed).  Row-major: segments are `pitch` doubles apart; tile-major: the G
// segments of a CTA are contiguous.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
__global__ void walk(const double *__restrict__ in, double *__restrict__ out, long tile_stride, long row_stride, int G, int depth) {
  const long base = (long)blockIdx.x * tile_stride + threadIdx.x * 2 - (threadIdx.x >= 64 ? 0 : 0);
  const double2 *src = reinterpret_cast<const double2 *>(in + (long)blockIdx.x * tile_stride) + (threadIdx.x & 63);
  double2 *dst = reinterpret_cast<double2 *>(out + (long)blockIdx.x * tile_stride) + (threadIdx.x & 63);
  (void)base;
  const int half = threadIdx.x >> 6;  // two half-CTAs take alternate rows: 64 threads x 16 B = one 1 KB segment
  double2 acc = make_double2(0, 0);
  for (int g = half; g < G; g += 2 * depth) {
    double2 v[8];
#pragma unroll
    for (int d = 0; d < 8; ++d)
      if (d < depth && g + 2 * d < G) v[d] = src[(long)(g + 2 * d) * row_stride / 2];
#pragma unroll
    for (int d = 0; d < 8; ++d)
      if (d < depth && g + 2 * d < G) {
        acc.x += v[d].x;
        dst[(long)(g + 2 * d) * row_stride / 2] = make_double2(v[d].x + 1.0, v[d].y + acc.x);
      }
  }
}
int main(int argc, char **argv) {
  const long N = argc > 1 ? atol(argv[1]) : 1000000;
  const int G = argc > 2 ? atoi(argv[2]) : 200;
  const long tiles = (N + 127) / 128, pitch = tiles * 128;
  double *a, *b;
  cudaMalloc(&a, sizeof(double) * pitch * G);
  cudaMalloc(&b, sizeof(double) * pitch * G);
  cudaMemset(a, 0, sizeof(double) * pitch * G);
  cudaMemset(b, 0, sizeof(double) * pitch * G);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int layout = 0; layout < 2; ++layout)
    for (int depth = 1; depth <= 8; depth *= 2) {
      const long ts = layout ? 128L * G : 128L, rs = layout ? 128L : pitch;
      for (int it = 0; it < 3; ++it) walk<<<tiles, 128>>>(a, b, ts, rs, G, depth);
      cudaEventRecord(e0);
      const int reps = 10;
      for (int it = 0; it < reps; ++it) walk<<<tiles, 128>>>(it & 1 ? b : a, it & 1 ? a : b, ts, rs, G, depth);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      printf("%s depth %d: %.3f ms/launch  %.0f GB/s (read+write)\n", layout ? "tile-major" : "row-major ", depth, ms / reps,
             2.0 * 8 * pitch * G / (ms / reps * 1e-3) / 1e9);
    }
  // sustained: the row-major walk, depth 4, back to back for about `seconds` (argv[3]) — what the same traffic reaches under
  // the board's power cap, which a 5 ms burst does not feel
  const double seconds = argc > 3 ? atof(argv[3]) : 0.0;
  if (seconds > 0.0) {
    const int reps = (int)(seconds / 0.5e-3);
    for (int chunk = 0; chunk < 4; ++chunk) {
      cudaEventRecord(e0);
      for (int it = 0; it < reps / 4; ++it) walk<<<tiles, 128>>>(it & 1 ? b : a, it & 1 ? a : b, 128L, pitch, G, 4);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      printf("sustained row-major depth 4, quarter %d: %.3f ms/launch  %.0f GB/s (read+write)\n", chunk, ms / (reps / 4),
             2.0 * 8 * pitch * G / (ms / (reps / 4) * 1e-3) / 1e9);
    }
  }
  printf("status %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
