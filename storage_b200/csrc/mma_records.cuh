// mma_records.cuh — per-step records of the FP64-mma sweep kernel (backward_sweep_mma.cu) and the code that folds them
// (called from fold_tables_kernel / reduce_fold_kernel in backward_sweep.cu and from the multi-GPU exchange kernel).
//
// One record per BLOCK of 8 inventory-grid points holds everything of the block that does not depend on the path, already
// in the order the kernel's lanes consume it (lane = 4 r + c: grid point r of the block, column c):
//   a[d][s][lane]   A fragments of the "payoff + continuation value" GEMM, mma.m8n8k4 row layout: coefficient 4 s + c of grid
//                   point r and decision d.  Coefficients 0..K-1 are the regression coefficients with the next-grid
//                   interpolation folded in (c = lw * gamma_lo + uw * gamma_hi, LSV.cs:282-305), coefficient K multiplies the
//                   spot price ((-v - consumed) * DF, LSV.cs:317-323), the rest are zero.
//   s[d][r]         offset = -(cost + inventory cost), the price coefficient again (for the realised value), and the weights
//                   of the two rows of V_{p+1} the decision lands between
//   rows[d][r]      byte offsets of those two rows in the kernel's shared-memory ring
// Grid points past the end of the period's grid (last block) repeat the last one; a grid point with fewer than three
// decisions repeats its first decision, which cannot win the strict > argmax (SH.cs:206-221).
#pragma once
#include <cstdint>

#include "device_types.hpp"
#include "kernels.hpp"

namespace lsmc {

struct MmaScalars {
  double offset, offset_again;  // twice: one 16-byte load fills the two-register start value of the mma accumulator pair
  double price_coef, lower_w, upper_w, pad;
};

template <int KSTEPS>
struct __align__(16) MmaBlockRecord {
  double a[3][KSTEPS][32];
  MmaScalars s[3][8];
  uint2 rows[3][8];
};

__host__ __device__ inline int mma_ksteps(int K) { return (K + 1 + 3) / 4; }  // basis functions + the price column, in steps of 4
__host__ __device__ inline size_t mma_record_bytes(int ksteps) { return (size_t)3 * ksteps * 32 * 8 + 3 * 8 * sizeof(MmaScalars) + 3 * 8 * sizeof(uint2); }

// item i = 3 * (padded grid point) + decision
__device__ __forceinline__ void fold_item_mma(const FoldArgs &f, int i) {
  const int gp = i / 3, d = i - 3 * gp;
  const int g = min(gp, f.G - 1);
  const int block = gp >> 3, r = gp & 7;
  const int ksteps = mma_ksteps(f.K);
  char *rec = static_cast<char *>(f.out) + (size_t)block * mma_record_bytes(ksteps);
  double *a = reinterpret_cast<double *>(rec) + (size_t)d * ksteps * 32;
  MmaScalars *sc = reinterpret_cast<MmaScalars *>(rec + (size_t)3 * ksteps * 32 * 8) + d * 8 + r;
  uint2 *rows = reinterpret_cast<uint2 *>(rec + (size_t)3 * ksteps * 32 * 8 + 3 * 8 * sizeof(MmaScalars)) + d * 8 + r;
  const int nd = f.num_decisions[g];
  const DevDecision e = f.decisions[(size_t)g * 3 + (d < nd ? d : 0)];
  const double price_coef = (-e.volume - e.consumed) * f.settle_df;
  const double offset = -(e.cost_npv + f.inventory_cost[g]);
  *sc = MmaScalars{offset, offset, price_coef, e.lower_w, e.upper_w, 0.0};
  *rows = make_uint2((uint32_t)(e.lower % f.ring_rows) * (uint32_t)f.row_pitch_bytes, (uint32_t)(e.upper % f.ring_rows) * (uint32_t)f.row_pitch_bytes);
  // gamma may have been written by other CTAs of the same launch: read it through L2
  const double *gl = f.gamma + (size_t)e.lower * f.K, *gu = f.gamma + (size_t)e.upper * f.K;
  for (int k = 0; k < 4 * ksteps; ++k) {
    double c = 0.0;
    if (k < f.K)
      c = (e.lower == e.upper) ? __ldcg(gl + k) : __ldcg(gu + k) * e.upper_w + __ldcg(gl + k) * e.lower_w;
    else if (k == f.K)
      c = price_coef;
    a[(size_t)(k >> 2) * 32 + r * 4 + (k & 3)] = c;
  }
}

__host__ __device__ inline int fold_item_count(const FoldArgs &f) { return f.format == 1 ? 3 * 8 * ((f.G + 7) / 8) : 3 * f.G; }

}  // namespace lsmc
