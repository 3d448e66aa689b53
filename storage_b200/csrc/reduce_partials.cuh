// reduce_partials.cuh — body of the per-step regression reduction, shared by reduce_partials_kernel (kernels.cu) and the
// kernel that also folds the next step's point tables in its last CTA (backward_sweep.cu).
#pragma once
#include <cuda_runtime.h>

#include "kernels.hpp"

namespace lsmc {

// gamma[g][k] = sum over chunks of partials[g][chunk][k], in a fixed order; one CTA per grid point.
// mean_mode: gamma[g][0] = sum / total_paths (the sample mean of LSV.cs:181), the other components are zero.
constexpr int kReduceThreads = 1024;  // stride = 1024 / (acc_k / 2) residue classes: a multiple of 16 for acc_k = 8 and 16
// smem_out != nullptr: the K sums are left in shared memory (first K entries) instead of a.out; the caller synchronises.
__device__ __forceinline__ void reduce_partials_body(const ReduceArgs &a, double *smem_out = nullptr) {
  const int g = blockIdx.x;
  // This grid point's partial sums are one contiguous run [num_chunks][acc_k].  A thread owns one pair of components
  // (one 16-byte load) of every `stride`-th chunk and keeps eight loads in flight; everything is combined in a fixed order.
  const int pairs = a.acc_k / 2;
  const int kp = threadIdx.x % pairs;
  const int lane_c = threadIdx.x / pairs;  // residue class of chunks
  const int stride = blockDim.x / pairs;
  const double2 *src = reinterpret_cast<const double2 *>(a.partials + (size_t)g * a.num_chunks * a.acc_k) + kp;
  double2 s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = make_double2(0.0, 0.0);
  // batches of eight independent loads per thread; the last, partial batch is predicated instead of walked one load at a
  // time (a run-time index into s[] would put the accumulators in local memory, and with fewer than 8 x stride chunks — any
  // rank of a sharded run — that tail is the whole reduction)
  for (int c = lane_c; c < a.num_chunks; c += 8 * stride) {
    double2 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int ci = c + i * stride;
      v[i] = ci < a.num_chunks ? src[(size_t)ci * pairs] : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      s[i].x += v[i].x;
      s[i].y += v[i].y;
    }
  }
  double2 sum;
  sum.x = ((s[0].x + s[1].x) + (s[2].x + s[3].x)) + ((s[4].x + s[5].x) + (s[6].x + s[7].x));
  sum.y = ((s[0].y + s[1].y) + (s[2].y + s[3].y)) + ((s[4].y + s[5].y) + (s[6].y + s[7].y));
  // combine the residue classes in a fixed order, in two levels: 16 runs of stride/16 classes, then the 16 run totals
  __shared__ double2 s_all[kReduceThreads];
  __shared__ double s_run[16][16];
  s_all[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x < 16 * a.acc_k) {
    const int k = threadIdx.x % a.acc_k, run = threadIdx.x / a.acc_k;
    const int per_run = stride / 16;
    double total = 0.0;
    for (int cc = run * per_run; cc < (run + 1) * per_run; ++cc) {
      const double2 v = s_all[cc * pairs + (k >> 1)];
      total += (k & 1) ? v.y : v.x;
    }
    s_run[run][k] = total;
  }
  __syncthreads();
  if (threadIdx.x < a.K) {
    double total = 0.0;
    for (int run = 0; run < 16; ++run) total += s_run[run][threadIdx.x];
    double out = total;
    if (a.mean_mode) out = (threadIdx.x == 0) ? total / (double)a.total_paths : 0.0;
    if (smem_out)
      smem_out[threadIdx.x] = out;
    else
      a.out[(size_t)g * a.K + threadIdx.x] = out;
  }
}


}  // namespace lsmc
