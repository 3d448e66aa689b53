// fp64_pipes.cu — microbenchmark behind the sweep kernel's design choice: does mma.sync.m8n8k4.f64 (DMMA) run on
// its own pipe beside DFMA on B200, and at what rate?  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_pipes fp64_pipes.cu
// Prints warp-instructions per clock per SM for DFMA-only, DMMA-only and an interleaved mix at several occupancies, and the
// dependent-issue latency of each.
#include <cuda_runtime.h>

#include <cstdio>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// MODE 0: DFMA only (CH independent chains); 1: DMMA only (CH independent accumulator pairs); 2: both, interleaved 1 DMMA : R DFMA
template <int MODE, int CH, int R>
__global__ void pipes(double *out, int iters, double a, double b, long long *cycles) {
  double f[CH * (MODE == 2 ? R : 1)], m0[CH], m1[CH];
#pragma unroll
  for (int i = 0; i < CH * (MODE == 2 ? R : 1); ++i) f[i] = threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < CH; ++i) m0[i] = m1[i] = threadIdx.x * 0.5 + i;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (MODE == 0) f[i] = __fma_rn(f[i], a, b);
      if (MODE == 1) dmma(m0[i], m1[i], a, b);
      if (MODE == 2) {
        dmma(m0[i], m1[i], a, b);
#pragma unroll
        for (int r = 0; r < R; ++r) f[i * R + r] = __fma_rn(f[i * R + r], a, b);
      }
    }
  }
  const long long t1 = clock64();
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < CH * (MODE == 2 ? R : 1); ++i) s += f[i];
#pragma unroll
  for (int i = 0; i < CH; ++i) s += m0[i] + m1[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE, int CH, int R>
void run(const char *name, int warps_per_sm, int sms, double *out, long long *d_cycles) {
  const int iters = 4096;
  const int threads = 32 * warps_per_sm;
  pipes<MODE, CH, R><<<sms, threads>>>(out, 16, 1.0000001, 1e-9, d_cycles);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  pipes<MODE, CH, R><<<sms, threads>>>(out, iters, 1.0000001, 1e-9, d_cycles);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  long long cyc = 0;
  cudaMemcpy(&cyc, d_cycles, sizeof(cyc), cudaMemcpyDeviceToHost);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double dfma = (MODE == 0 ? CH : (MODE == 2 ? CH * R : 0)) * (double)iters * warps_per_sm;
  const double dm = (MODE >= 1 ? CH : 0) * (double)iters * warps_per_sm;
  std::printf("%-28s warps/SM %2d  cycles %9lld  DFMA/clk/SM %.3f  DMMA/clk/SM %.4f  (%.3f ms)%s\n", name, warps_per_sm, cyc, dfma / cyc, dm / cyc, ms,
              cudaGetLastError() == cudaSuccess ? "" : "  CUDA ERROR");
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double *out;
  long long *cyc;
  cudaMalloc(&out, sizeof(double) * sms * 1024);
  cudaMalloc(&cyc, sizeof(long long));
  std::printf("latency (1 warp per SM, 1 chain):\n");
  run<0, 1, 1>("DFMA dependent chain", 1, sms, out, cyc);
  run<1, 1, 1>("DMMA dependent chain", 1, sms, out, cyc);
  std::printf("throughput:\n");
  for (int w : {4, 8, 12, 16, 32}) {
    run<0, 8, 1>("DFMA x8 chains", w, sms, out, cyc);
    run<1, 8, 1>("DMMA x8 chains", w, sms, out, cyc);
    run<2, 4, 2>("1 DMMA : 2 DFMA", w, sms, out, cyc);
    run<2, 4, 4>("1 DMMA : 4 DFMA", w, sms, out, cyc);
    run<2, 2, 8>("1 DMMA : 8 DFMA", w, sms, out, cyc);
  }
  return 0;
}
