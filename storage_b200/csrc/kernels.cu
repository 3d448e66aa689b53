// kernels.cu — hand-written sm_100a kernels of the LSMC hot path.
//
//   gram_kernel / cholesky_kernel   regression pre-pass: double-double Gram matrix of the monomial design matrix of every
//                                   period (LSV.cs:838-855 PopulateDesignMatrix), its Cholesky factor R and R^-1, i.e. the
//                                   thin-QR of LSV.cs:193-196 formed without ever materialising Q.
//   basis_kernel                    orthonormal basis rows Q = X R^-1 of one period's regression simulations, materialised.
//   backward_step_kernel            the GENERAL backward-induction step (LSV.cs:240-338): any decision-table width
//                                   (ExtraDecisions > 0) and the terminal step.  The usual three-decision tables run on
//                                   the sweep kernels of backward_sweep.cu.  Continuation value = q(path) . c(grid point,
//                                   decision), decision argmax, storage value; fused with the accumulation of q_prev^T V
//                                   for the regression of the next backward step, so every storage value is written once
//                                   and read once.
//   reduce_partials_kernel          fixed-order reduction of the per-tile regression partial sums -> gamma (LSV.cs:204).
//   forward_kernel                  the whole forward simulation (LSV.cs:390-521) in one launch: periods outer, a CTA's path
//                                   tiles inner, path state in shared memory; per-period sums reduced per CTA in a fixed order.
//   sum kernels                     deterministic reductions for NPV / standard error / backward NPV.
//
// HBM-bound FP64 streaming work; the only tensor-core use is the skinny FP64 GEMM of the regression sums
// (mma.sync m8n8k4 f64).  Paths are the contiguous dimension of every buffer so warps read and write whole 128-byte lines;
// the storage-value slabs use a row pitch that is a multiple of 256 paths and are accessed as double2.
#include <cuda_runtime.h>

#include <cstdint>
#include <algorithm>
#include <cstdlib>

#include "backward_common.cuh"
#include "dd.cuh"
#include "device_types.hpp"
#include "kernels.hpp"
#include "reduce_partials.cuh"

#ifndef LSMC_GRAM_MIN_CTAS
#define LSMC_GRAM_MIN_CTAS 2  // 252 registers; 3 CTAs (168 registers) measured equal, 4 CTAs spill and are 4x slower
#endif
#ifndef LSMC_FWD_MIN_CTAS
#define LSMC_FWD_MIN_CTAS 4  // 124 registers: measured 23 % faster than the 164-register build (3 CTAs/SM)
#endif

namespace lsmc {

namespace {

constexpr unsigned kFullMask = 0xffffffffu;

// ------------------------------------------------------------------------------------------------------------------
// Orthonormalised basis row of one path: monomials X_k (BasisFunctions.Generic, BasisFunctions.cs:93-135; integer powers
// correctly rounded, products in factor-index order) followed by q_j = sum_{l<=j} X_l * Rinv[l][j] accumulated with
// compensated (twice-the-working-precision) summation.  Loops over factors are deliberately not unrolled: this code is
// cold relative to the decision loop and must stay small in the instruction cache.
// ------------------------------------------------------------------------------------------------------------------
// The exponent table compiled into a term list once per CTA: monomial k is the product of var[t]^pow[t] for
// t in [begin[k], begin[k+1]), spot first, then factors in index order (only the non-zero exponents).
constexpr int kMaxTerms = LSMC_MAX_BASIS * (1 + LSMC_MAX_FACTORS);
struct MonomialProgram {
  int32_t begin[LSMC_MAX_BASIS + 1];
  int8_t var[kMaxTerms];
  int8_t pow[kMaxTerms];
};
__device__ __forceinline__ void build_monomial_program(MonomialProgram &pr, const int32_t *__restrict__ pw, int K, int F) {  // one thread
  int n = 0;
  for (int k = 0; k < K; ++k) {
    pr.begin[k] = n;
    for (int v = 0; v <= F; ++v) {
      const int p = pw[k * (1 + F) + v];
      if (p > 0) {
        pr.var[n] = (int8_t)v;
        pr.pow[n] = (int8_t)p;
        ++n;
      }
    }
  }
  for (int k = K; k <= LSMC_MAX_BASIS; ++k) pr.begin[k] = n;
}

// The simulated inputs of one path (spot and factors of one period), loaded one loop iteration ahead of their use so the
// global-load latency hides behind the previous path's arithmetic.
struct PathInputs {
  double s;
  double x[LSMC_MAX_FACTORS];
};
__device__ __forceinline__ void load_inputs(const BasisSpec &b, int F, int64_t n, PathInputs &in) {
  in.s = b.spot[n];
#pragma unroll
  for (int i = 0; i < LSMC_MAX_FACTORS; ++i)
    if (i < F) in.x[i] = b.factors[i][n];
}
template <int KP>
__device__ __forceinline__ void eval_monomials_from(const PathInputs &in, const MonomialProgram &pr, int K, int F, double (&X)[KP]) {
  double v[1 + LSMC_MAX_FACTORS];  // indexed by the program: lives in local memory (L1)
  v[0] = in.s;
#pragma unroll
  for (int i = 0; i < LSMC_MAX_FACTORS; ++i)
    if (i < F) v[1 + i] = in.x[i];
#pragma unroll
  for (int k = 0; k < KP; ++k) {
    double m = 1.0;  // 1.0 * x is exact: the first factor enters unrounded, as in the reference's product
    if (k < K) {
      const int end = pr.begin[k + 1];
#pragma unroll 1
      for (int t = pr.begin[k]; t < end; ++t) m = m * pow_int_rounded(v[pr.var[t]], pr.pow[t]);
    }
    X[k] = (k < K) ? m : 0.0;
  }
}

// q_j = sum_{l<=j} X_l * (R.hi + R.lo)[l][j]: error-free products, cascaded (Dot2) summation, one final rounding
template <int KP>
__device__ __forceinline__ void orthonormalise(const double (&X)[KP], const dd *__restrict__ rinv, int K, double (&q)[KP]) {
#pragma unroll
  for (int j = 0; j < KP; ++j) {
    double acc = 0.0, comp = 0.0;
    if (j < K) {
#pragma unroll
      for (int l = 0; l <= j; ++l) {
        const dd r = rinv[l * K + j];
        const double p = X[l] * r.hi;
        const double e = __fma_rn(X[l], r.hi, -p);
        const double s = acc + p;
        const double bb = s - acc;
        comp += ((acc - (s - bb)) + (p - bb)) + __fma_rn(X[l], r.lo, e);
        acc = s;
      }
    }
    q[j] = acc + comp;
  }
}

template <int KP>
__device__ __forceinline__ void eval_q(const BasisSpec &b, const MonomialProgram &pr, const dd *__restrict__ rinv, int K,
                                       int F, int64_t n, bool valid, double (&q)[KP]) {
#pragma unroll
  for (int k = 0; k < KP; ++k) q[k] = 0.0;
  if (!valid) return;
  if (b.mean_mode) {
    q[0] = 1.0;
    return;
  }
  double X[KP];
  PathInputs in;
  load_inputs(b, F, n, in);
  eval_monomials_from<KP>(in, pr, K, F, X);
  orthonormalise<KP>(X, rinv, K, q);
}

// Sum of 8 per-lane values over the warp with 9 shuffles: after the call every lane holds the warp total of component
// kidx = 4*bit4(lane) + 2*bit3(lane) + bit2(lane).  The pairing order is fixed, so results are reproducible.
__device__ __forceinline__ double warp_reduce8(const double (&v)[8], int lane, int &kidx) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
  double w[4], u[2];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double send = b4 ? v[i] : v[i + 4];
    double keep = b4 ? v[i + 4] : v[i];
    w[i] = keep + __shfl_xor_sync(kFullMask, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    double send = b3 ? w[i] : w[i + 2];
    double keep = b3 ? w[i + 2] : w[i];
    u[i] = keep + __shfl_xor_sync(kFullMask, send, 8);
  }
  double send = b2 ? u[0] : u[1];
  double keep = b2 ? u[1] : u[0];
  double t = keep + __shfl_xor_sync(kFullMask, send, 4);
  t += __shfl_xor_sync(kFullMask, t, 2);
  t += __shfl_xor_sync(kFullMask, t, 1);
  kidx = (b4 ? 4 : 0) + (b3 ? 2 : 0) + (b2 ? 1 : 0);
  return t;
}

__device__ __forceinline__ dd warp_reduce_dd(dd v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    dd o;
    o.hi = __shfl_xor_sync(kFullMask, v.hi, off);
    o.lo = __shfl_xor_sync(kFullMask, v.lo, off);
    v = dd_add(v, o);
  }
  return v;
}

// ------------------------------------------------------------------------------------------------------------------
// Regression pre-pass
// ------------------------------------------------------------------------------------------------------------------
// grid = (chunks, periods); each CTA accumulates the upper triangle of X^T X over its path chunk in double-double.
template <int KP>
__global__ void __launch_bounds__(kThreads, LSMC_GRAM_MIN_CTAS) gram_kernel(GramArgs a) {
  constexpr int NP = KP * (KP + 1) / 2;
  __shared__ int32_t s_pw[LSMC_MAX_BASIS * (1 + LSMC_MAX_FACTORS)];
  __shared__ dd s_red[kWarps][NP];
  const int slot = blockIdx.y;
  const BasisSpec b = a.specs[slot];
  const int K = a.K, F = a.F;
  __shared__ MonomialProgram s_prog;
  for (int i = threadIdx.x; i < K * (1 + F); i += blockDim.x) s_pw[i] = a.powers[i];
  __syncthreads();
  if (threadIdx.x == 0) build_monomial_program(s_prog, s_pw, K, F);
  __syncthreads();
  // compensated accumulation: hi carries the running sum, lo collects the rounding errors of the sums and products
  double hi[NP], lo[NP];
#pragma unroll
  for (int i = 0; i < NP; ++i) hi[i] = lo[i] = 0.0;
  const int64_t begin = (int64_t)blockIdx.x * a.chunk_paths;
  const int64_t end = min(begin + a.chunk_paths, a.N);
  PathInputs next;
  if (begin + threadIdx.x < end) load_inputs(b, F, begin + threadIdx.x, next);
  for (int64_t n = begin + threadIdx.x; n < end; n += blockDim.x) {
    const PathInputs in = next;
    if (n + blockDim.x < end) load_inputs(b, F, n + blockDim.x, next);
    double X[KP];
    eval_monomials_from<KP>(in, s_prog, K, F, X);
    int idx = 0;
#pragma unroll
    for (int i = 0; i < KP; ++i)
#pragma unroll
      for (int j = i; j < KP; ++j, ++idx) {
        const double p = X[i] * X[j];
        const double e = __fma_rn(X[i], X[j], -p);
        const double s = hi[idx] + p;
        const double bb = s - hi[idx];
        lo[idx] += ((hi[idx] - (s - bb)) + (p - bb)) + e;
        hi[idx] = s;
      }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NP; ++i) {
    dd r = warp_reduce_dd(quick_two_sum(hi[i], lo[i]));
    if (lane == 0) s_red[warp][i] = r;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NP; i += blockDim.x) {
    dd r = s_red[0][i];
    for (int w = 1; w < kWarps; ++w) r = dd_add(r, s_red[w][i]);
    double *dst = a.partials + (((size_t)slot * gridDim.x + blockIdx.x) * NP + i) * 2;
    dst[0] = r.hi;
    dst[1] = r.lo;
  }
}

// One CTA per regression period: fixed-order sum of the chunk partials (and, when sharded, of the other ranks' sums that
// the host gathered into `partials`), Cholesky A = R^T R and R^-1, all in double-double.  Writes rinv[slot][K*K] as
// (hi,lo) pairs, row-major [l][j], zero below the diagonal.
template <int KP>
__global__ void cholesky_kernel(CholeskyArgs a) {
  constexpr int NP = KP * (KP + 1) / 2;
  __shared__ dd s_gram[NP];
  const int slot = blockIdx.x;
  const int K = a.K;
  for (int i = threadIdx.x; i < NP; i += blockDim.x) {
    dd r{0.0, 0.0};
    for (int rk = 0; rk < a.num_ranks; ++rk)
      for (int c = 0; c < a.num_chunks; ++c) {
        const double *src = a.partials + ((((size_t)rk * a.num_slots + slot) * a.num_chunks + c) * NP + i) * 2;
        r = dd_add(r, dd{src[0], src[1]});
      }
    s_gram[i] = r;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  dd R[KP][KP], Rinv[KP][KP];
  auto gram = [&](int i, int j) {  // i <= j, packed upper triangle of a KP x KP matrix
    return s_gram[i * KP - i * (i - 1) / 2 + (j - i)];
  };
  bool ok = true;
  for (int j = 0; j < K; ++j) {
    for (int i = 0; i <= j; ++i) {
      dd sum = gram(i, j);
      for (int l = 0; l < i; ++l) sum = dd_sub(sum, dd_mul(R[l][i], R[l][j]));
      if (i == j) {
        if (!(sum.hi > 0.0)) {
          ok = false;
          sum = dd{1.0, 0.0};
        }
        R[i][j] = dd_sqrt(sum);
      } else {
        R[i][j] = dd_div(sum, R[i][i]);
      }
    }
  }
  // A pivot that lost all its digits means the monomials are linearly dependent on the simulated paths.  The Gram matrix
  // carries ~32 digits (compensated accumulation, double-double factorisation), so pivots down to 1e-28 of the diagonal are
  // still resolved; nearly collinear but independent columns (e.g. powers of a spot price that barely moves) are legitimate
  // input that the reference's QR accepts as well.
  for (int j = 0; j < K; ++j) {
    dd diag = gram(j, j);
    double rel = (R[j][j].hi * R[j][j].hi) / diag.hi;
    if (!(rel > 1e-28)) ok = false;
  }
  if (!ok) atomicExch(a.error_flag, (int)LSMC_ERR_NUMERICAL);
  for (int j = 0; j < K; ++j)
    for (int i = j; i >= 0; --i) {
      dd sum = (i == j) ? dd{1.0, 0.0} : dd{0.0, 0.0};
      for (int l = i + 1; l <= j; ++l) sum = dd_sub(sum, dd_mul(R[i][l], Rinv[l][j]));
      Rinv[i][j] = dd_div(sum, R[i][i]);
    }
  double *out = a.rinv + (size_t)slot * K * K * 2;
  for (int l = 0; l < K; ++l)
    for (int j = 0; j < K; ++j) {
      dd v = (l <= j) ? Rinv[l][j] : dd{0.0, 0.0};
      out[(l * K + j) * 2] = v.hi;
      out[(l * K + j) * 2 + 1] = v.lo;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// Orthonormal basis rows of one period, materialised once: Q[k][n] for the regression simulations.  32 B read and 8K B
// written per path; the backward kernel then reads the rows of period p (continuation values) and p-1 (regression sums)
// as coalesced double2 instead of re-deriving them per inventory-grid tile.
// ------------------------------------------------------------------------------------------------------------------
template <int KP>
__global__ void __launch_bounds__(256) basis_kernel(BasisArgs a) {
  __shared__ int32_t s_pw[LSMC_MAX_BASIS * (1 + LSMC_MAX_FACTORS)];
  __shared__ dd s_rinv[LSMC_MAX_BASIS * LSMC_MAX_BASIS];
  const int K = a.K, F = a.F;
  __shared__ MonomialProgram s_prog;
  for (int i = threadIdx.x; i < K * (1 + F); i += blockDim.x) s_pw[i] = a.powers[i];
  for (int i = threadIdx.x; i < K * K; i += blockDim.x) s_rinv[i] = dd{a.spec.rinv[2 * i], a.spec.rinv[2 * i + 1]};
  __syncthreads();
  if (threadIdx.x == 0) build_monomial_program(s_prog, s_pw, K, F);
  __syncthreads();
  // a CTA walks 256-path chunks with a grid stride: the table staging above (two barriers, one thread compiling the exponent
  // table) is paid once per CTA instead of once per 256 paths (it was 27 % of this kernel's stall samples)
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < a.pitch; n += (int64_t)gridDim.x * blockDim.x) {
    if (n >= a.N) {  // padding columns of the rows stay zero: the backward kernels load them as the odd half of a double2
      for (int k = 0; k < K; ++k) a.q_out[(size_t)k * a.pitch + n] = 0.0;
      continue;
    }
    double q[KP];
    eval_q<KP>(a.spec, s_prog, s_rinv, K, F, n, true, q);
#pragma unroll
    for (int k = 0; k < KP; ++k)
      if (k < K) a.q_out[(size_t)k * a.pitch + n] = q[k];
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Backward induction step
// ------------------------------------------------------------------------------------------------------------------
// One CTA = one tile of 256 paths (128 threads x double2) x a contiguous range of inventory-grid sub-tiles.  Within the
// CTA the grid points are swept in ascending order, so the rows of v_next a decision interpolates between are re-used
// from L1 by the neighbouring grid points and each row comes from HBM once.
//
// The regression sums of the next backward step, sum_n q_prev[n][k] * V[g][n], are a dense skinny FP64 GEMM
// (K x paths) x (paths x grid points).  Each warp runs it on the tensor cores (mma.sync m8n8k4 f64) over its own 64
// paths from a warp-private shared-memory tile, 8 grid points at a time; warps are combined in a fixed order.
#ifndef LSMC_PF_DIST
#define LSMC_PF_DIST 4
#endif
#ifndef LSMC_PF_L1
#define LSMC_PF_L1 0
#endif
#ifndef LSMC_BACKWARD_MIN_CTAS
#define LSMC_BACKWARD_MIN_CTAS 4
#endif
// DFIX: decision-table row width when it is the usual 3 (no extra decisions) so the table indexing folds at compile time;
// 0 = run-time width.
// FUSED (only with DFIX == 3): fused-payoff arithmetic of backward_common.cuh's point_fused instead of the reference's
// operation order.
template <int KP, int DFIX, bool FUSED>
__global__ void __launch_bounds__(kThreads, LSMC_BACKWARD_MIN_CTAS) backward_step_kernel(BackwardStepArgs a) {
  constexpr int ACC = KP <= 8 ? 8 : 16;       // width of a stored partial-sum row
  constexpr int KC = (KP + 1) & ~1;           // coefficient row padded to an even length (16-byte rows)
  constexpr int TG = kMaxGridTile;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int K = a.K, D = DFIX ? DFIX : a.D;
  // dynamic smem: decA [TG][D] | decB [TG][D] | coef [TG][D][KC] | vtile [kWarps][8][pitch] | qtile [kWarps][ACC][pitch]
  //               | acc [kWarps][ACC][TG] | inv_cost [TG] | nd [TG]
  DecisionA *s_a = reinterpret_cast<DecisionA *>(smem_raw);
  DecisionR *s_b = reinterpret_cast<DecisionR *>(s_a + (size_t)TG * D);
  double *s_coef = reinterpret_cast<double *>(s_b + (size_t)TG * D);
  double *s_vtile = s_coef + (size_t)TG * D * KC;
  double *s_qtile = s_vtile + (size_t)kWarps * 8 * kTilePitch;
  double *s_acc = s_qtile + (size_t)kWarps * ACC * kTilePitch;
  double *s_inv_cost = s_acc + (size_t)kWarps * ACC * TG;
  int32_t *s_nd = reinterpret_cast<int32_t *>(s_inv_cost + TG);

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.y;
  const int64_t n0 = tile * kTilePaths + (int64_t)threadIdx.x * kPathsPerThread;
  const bool valid0 = n0 < a.N, valid1 = n0 + 1 < a.N;
  const int64_t c0 = valid0 ? n0 : ((a.N - 1) & ~int64_t(1));  // clamped, even: vector loads stay inside the padded row
  const int64_t c1 = valid1 ? n0 + 1 : a.N - 1;

  double price0, price1;
  if (a.spot) {
    price0 = a.spot[valid0 ? n0 : a.N - 1];
    price1 = a.spot[c1];
  } else {
    price0 = price1 = a.const_price;
  }
  // basis row of this period (continuation value)
  double q0[KP], q1[KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) {
    q0[k] = q1[k] = 0.0;
    if (k < K && a.q_this) {
      const double2 t = *reinterpret_cast<const double2 *>(a.q_this + (size_t)k * a.pitch + c0);
      q0[k] = t.x;
      q1[k] = t.y;
    }
  }
  if (!a.terminal_mode && !a.q_this) q0[0] = q1[0] = 1.0;  // mean mode: continuation value = gamma[g][0]

  // basis rows of the previous period for the regression sums, staged as the A operand of the warp's GEMM
  double *my_q = s_qtile + (size_t)warp * ACC * kTilePitch;
  double *my_v = s_vtile + (size_t)warp * 8 * kTilePitch;
  if (a.accumulate) {
#pragma unroll
    for (int k = 0; k < ACC; ++k) {
      double2 t = make_double2(0.0, 0.0);
      if (k < K) {
        if (a.q_prev) {
          t = *reinterpret_cast<const double2 *>(a.q_prev + (size_t)k * a.pitch + c0);
        } else if (k == 0) {
          t = make_double2(1.0, 1.0);  // next regression is the sample mean: plain sum of V
        }
      }
      if (!valid0) t.x = 0.0;
      if (!valid1) t.y = 0.0;
      *reinterpret_cast<double2 *>(my_q + (size_t)k * kTilePitch + 2 * lane) = t;
    }
  }
  const double df = a.settle_df;
  const int sub_begin = blockIdx.x * a.subtiles_per_cta;
  const int sub_end = min(sub_begin + a.subtiles_per_cta, (a.G + TG - 1) / TG);
  const double *vbase = a.v_next + c0;  // this thread's two columns of every v_next row
  const uint32_t last_row_off = (uint32_t)(a.Gn > 0 ? a.Gn - 1 : 0) * (uint32_t)a.pitch;

  for (int sub = sub_begin; sub < sub_end; ++sub) {
    const int g0 = sub * TG;
    const int gcount = min(TG, a.G - g0);
    __syncthreads();  // the previous sub-tile's tables and partial sums are no longer in use
    // ---- stage this sub-tile's sim-independent tables; fold the next-grid interpolation into the coefficients
    if (!a.terminal_mode) {
      for (int i = threadIdx.x; i < gcount * D; i += blockDim.x) {
        const int g = i / D;
        if (i - g * D < a.num_decisions[g0 + g]) {
          const DevDecision e = a.decisions[(size_t)g0 * D + i];
          if (FUSED)
            s_a[i] = DecisionA{(-e.volume - e.consumed) * a.settle_df, -(e.cost_npv + a.inventory_cost[g0 + g]), e.lower_w, e.upper_w};
          else
            s_a[i] = DecisionA{-e.volume, -e.consumed, e.cost_npv, e.lower_w};
          s_b[i] = DecisionR{e.upper_w, (uint32_t)e.lower * (uint32_t)a.pitch, (uint32_t)e.upper * (uint32_t)a.pitch};
          fold_coefficients<KC>(e, a.gamma, K, s_coef + (size_t)i * KC);
        }
      }
      for (int i = threadIdx.x; i < gcount; i += blockDim.x) {
        s_inv_cost[i] = a.inventory_cost[g0 + i];
        s_nd[i] = a.num_decisions[g0 + i];
      }
    }
    __syncthreads();

    // ---- sweep the sub-tile's grid points.  Everything that varies with g advances by pointer increments; the rows of
    // v_next for grid point g+1 are requested before the arithmetic of grid point g (two row buffers used alternately).
    const DecisionA *pa = s_a;
    const DecisionR *pb = s_b;
    const double *pc = s_coef;
    double *vout = a.v_this + (size_t)g0 * a.pitch + n0;
    Rows rows_a, rows_b;
    const bool fast = !a.terminal_mode && DFIX == 3;
    if (fast) load_rows(pb, s_nd[0], vbase, rows_a);
    for (int g = 0; g < gcount; g += 2) {
      double v0, v1, w0 = 0.0, w1 = 0.0;
      const bool second = g + 1 < gcount;
      if (fast) {
#if LSMC_PF_DIST > 0
        {
          // The row window slides up about one row per grid point; the rows a few grid points ahead are still in HBM.
          // Ask for them now so they are in cache when their turn comes (one 16-byte column per thread = whole lines per warp).
          const uint32_t ahead = pb[s_nd[g] - 1].upper_off + (uint32_t)LSMC_PF_DIST * (uint32_t)a.pitch;
          const double *pf = vbase + (ahead < last_row_off ? ahead : last_row_off);
#if LSMC_PF_L1
          asm volatile("prefetch.global.L1 [%0];" ::"l"(pf));
          asm volatile("prefetch.global.L1 [%0];" ::"l"(pf + a.pitch < vbase + last_row_off ? pf + a.pitch : vbase + last_row_off));
#else
          asm volatile("prefetch.global.L2 [%0];" ::"l"(pf));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(pf + a.pitch < vbase + last_row_off ? pf + a.pitch : vbase + last_row_off));
#endif
        }
#endif
        if (second) load_rows(pb + 3, s_nd[g + 1], vbase, rows_b);
        if (FUSED)
          point_fused<KP, KC>(reinterpret_cast<const DecisionF *>(pa), pc, s_nd[g], rows_a, q0, q1, price0, price1, v0, v1);
        else
          point<KP, KC>(pa, pb, pc, s_nd[g], s_inv_cost[g], rows_a, q0, q1, price0, price1, df, v0, v1);
        if (g + 2 < gcount) load_rows(pb + 6, s_nd[g + 2], vbase, rows_a);
        if (second) {
          if (FUSED)
            point_fused<KP, KC>(reinterpret_cast<const DecisionF *>(pa + 3), pc + 3 * KC, s_nd[g + 1], rows_b, q0, q1, price0, price1, w0, w1);
          else
            point<KP, KC>(pa + 3, pb + 3, pc + 3 * KC, s_nd[g + 1], s_inv_cost[g + 1], rows_b, q0, q1, price0, price1, df, w0, w1);
        }
      } else if (a.terminal_mode) {
        // V_T[g][n] = TerminalStorageNpv(S_T[n], grid_T[g]) (LSV.cs:121-131)
        const double inv = a.grid_this[g0 + g], inv2 = second ? a.grid_this[g0 + g + 1] : 0.0;
        v0 = ((a.terminal[3] * (price0 * inv) + a.terminal[1] * price0) + a.terminal[2] * inv) + a.terminal[0];
        v1 = ((a.terminal[3] * (price1 * inv) + a.terminal[1] * price1) + a.terminal[2] * inv) + a.terminal[0];
        w0 = ((a.terminal[3] * (price0 * inv2) + a.terminal[1] * price0) + a.terminal[2] * inv2) + a.terminal[0];
        w1 = ((a.terminal[3] * (price1 * inv2) + a.terminal[1] * price1) + a.terminal[2] * inv2) + a.terminal[0];
      } else {
        point_generic<KP, KC>(s_a + g * D, s_b + g * D, s_coef + (size_t)g * D * KC, s_nd[g], s_inv_cost[g], vbase, q0, q1, price0,
                              price1, df, v0, v1);
        if (second)
          point_generic<KP, KC>(s_a + (g + 1) * D, s_b + (g + 1) * D, s_coef + (size_t)(g + 1) * D * KC, s_nd[g + 1], s_inv_cost[g + 1],
                                vbase, q0, q1, price0, price1, df, w0, w1);
      }
      if (!valid0) v0 = w0 = 0.0;
      if (!valid1) v1 = w1 = 0.0;
      if (valid0) {
        *reinterpret_cast<double2 *>(vout) = make_double2(v0, v1);
        if (second) *reinterpret_cast<double2 *>(vout + a.pitch) = make_double2(w0, w1);
      }
      vout += 2 * a.pitch;
      pa += 6;
      pb += 6;
      pc += 6 * KC;
      if (a.accumulate) {
        double *vt = my_v + (size_t)(g & 7) * kTilePitch + 2 * lane;
        *reinterpret_cast<double2 *>(vt) = make_double2(v0, v1);
        *reinterpret_cast<double2 *>(vt + kTilePitch) = make_double2(w0, w1);
        if ((g & 7) == 6 || g + 2 >= gcount) {
          // C[k][g] = sum over the warp's 64 paths of Q_prev[k][n] * V[g][n] for this group of 8 grid points (rows of the
          // tile beyond the last grid point hold stale values; their columns of C are never read)
          __syncwarp();
          const int row = lane >> 2, col = lane & 3, gg = g & ~7;
#pragma unroll
          for (int mb = 0; mb < ACC; mb += 8) {
            double acc0 = 0.0, acc1 = 0.0;
            const double *qa = my_q + (size_t)(mb + row) * kTilePitch + col;
            const double *vb = my_v + (size_t)row * kTilePitch + col;
#pragma unroll 4
            for (int p = 0; p < kWarpPaths; p += 4) dmma_m8n8k4(acc0, acc1, qa[p], vb[p]);
            // C fragment: row = basis index mb + lane/4, columns = grid points 2*(lane%4) + {0,1} of this group
            double *dst = s_acc + ((size_t)warp * ACC + mb + row) * TG + gg + 2 * col;
            dst[0] = acc0;
            dst[1] = acc1;
          }
          __syncwarp();
        }
      }
    }

    if (a.accumulate) {
      __syncthreads();
      for (int i = threadIdx.x; i < gcount * ACC; i += blockDim.x) {
        const int g = i / ACC, k = i - g * ACC;
        double sum = s_acc[((size_t)0 * ACC + k) * TG + g];
        for (int w = 1; w < kWarps; ++w) sum += s_acc[((size_t)w * ACC + k) * TG + g];
        a.partials[((size_t)(g0 + g) * gridDim.y + tile) * ACC + k] = sum;
      }
    }
  }
}

// gamma[g][k] = sum over chunks of partials[g][chunk][k], in a fixed order; one CTA per grid point.
// mean_mode: gamma[g][0] = sum / total_paths (the sample mean of LSV.cs:181), the other components are zero.
__global__ void __launch_bounds__(kReduceThreads) reduce_partials_kernel(ReduceArgs a) { reduce_partials_body(a); }

// ------------------------------------------------------------------------------------------------------------------
// Forward simulation
// ------------------------------------------------------------------------------------------------------------------
struct DevRange {
  double lo, hi;
};

__device__ __forceinline__ int left_segment_dev(const double *x, int n, double t) {
  int lo = 0, hi = n;  // first index with x[idx] > t
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (x[mid] > t)
      hi = mid;
    else
      lo = mid + 1;
  }
  return lo - 1;  // last knot <= t (-1 if none)
}

__device__ __forceinline__ double spline_dev(const double *x, const double *y, int n, double t) {
  int k = left_segment_dev(x, n, t);
  k = max(0, min(k, n - 2));
  const double slope = (y[k + 1] - y[k]) / (x[k + 1] - x[k]);
  return y[k] + (t - x[k]) * slope;
}

__device__ __forceinline__ double horner_dev(const double *c, int len, double z) {
  double acc = c[len - 1];
  for (int i = len - 2; i >= 0; --i) {
    acc *= z;
    acc += c[i];
  }
  return acc;
}

// CmdtyStorage.GetInjectWithdrawRange (CS.cs:86-100) at a per-path inventory
__device__ __forceinline__ DevRange storage_range_dev(const ForwardArgs &a, const DevPeriodParams &pp, double inventory, int *err) {
  if (inventory < pp.min_inventory || inventory > pp.max_inventory) *err = LSMC_ERR_ARGUMENT;
  const double *xi = a.ratchet_inventory + pp.rbegin, *mn = a.ratchet_min + pp.rbegin, *mx = a.ratchet_max + pp.rbegin;
  switch (pp.kind) {
    case LSMC_CONSTRAINT_CONSTANT:
      return {mn[0], mx[0]};
    case LSMC_CONSTRAINT_PIECEWISE_LINEAR: {
      // MathNet LinearSpline: y_k + (t - x_k) * slope_k with the slopes formed once (host, same division), one segment
      // search for both splines
      int k = left_segment_dev(xi, pp.rcount, inventory);
      k = max(0, min(k, pp.rcount - 2));
      const double dx = inventory - xi[k];
      return {mn[k] + dx * a.ratchet_min_slope[pp.rbegin + k], mx[k] + dx * a.ratchet_max_slope[pp.rbegin + k]};
    }
    case LSMC_CONSTRAINT_STEP: {
      if (inventory < xi[0] || inventory > xi[pp.rcount - 1]) *err = LSMC_ERR_ARGUMENT;
      int idx = max(0, min(left_segment_dev(xi, pp.rcount, inventory), pp.rcount - 1));
      return {mn[idx], mx[idx]};
    }
    default:
      return {horner_dev(a.poly_min + pp.poly_begin, pp.rcount, inventory), horner_dev(a.poly_max + pp.poly_begin, pp.rcount, inventory)};
  }
}

// StorageHelper.CalculateBangBangDecisionSet (SH.cs:109-204)
__device__ __forceinline__ int decision_set_dev(DevRange r, double inventory, double loss, double next_min, double next_max, double tol,
                                                int extra, double *out, int *err) {
  const double base = inventory - loss;
  double withdraw, inject;
  const double lowest = r.lo + base;
  if (lowest > next_max) {
    if (!(lowest - next_max < tol)) *err = LSMC_ERR_ARGUMENT;
    withdraw = next_max - base;
  } else {
    withdraw = lowest > next_min ? r.lo : next_min - base;
  }
  const double highest = r.hi + base;
  if (highest < next_min) {
    if (!(next_min - highest < tol)) *err = LSMC_ERR_ARGUMENT;
    inject = next_min - base;
  } else {
    inject = highest < next_max ? r.hi : next_max - base;
  }
  int n = 0;
  out[n++] = withdraw;
  if (withdraw >= 0.0 || inject <= 0.0) {
    const double step = (inject - withdraw) / (extra + 1);
    for (int i = 0; i < extra; ++i) out[n++] = (i + 1) * step + withdraw;
  } else {
    const double step_w = (0 - withdraw) / (extra + 1);
    for (int i = 0; i < extra; ++i) out[n++] = (i + 1) * step_w + withdraw;
    out[n++] = 0.0;
    const double step_i = (inject - 0) / (extra + 1);
    for (int i = 0; i < extra; ++i) out[n++] = (i + 1) * step_i + 0;
  }
  out[n++] = inject;
  return n;
}

// StorageHelper.BisectInventorySpace (SH.cs:280-314)
__device__ __noinline__ void bisect_dev(const double *grid, int n, double inventory, double tol, int *lower, int *upper, int *err) {
  if (n == 1 && fabs(inventory - grid[0]) <= tol) {
    *lower = *upper = 0;
    return;
  }
  int left = 0, right = n - 1;
  const int top = right;
  while (right > left) {
    const int mid = (left + right) / 2;
    const double gm = grid[mid];
    if (fabs(inventory - gm) <= tol) {
      *lower = *upper = mid;
      return;
    }
    if (gm > inventory) {
      right = mid;
      continue;
    }
    const double g1 = grid[mid + 1];
    if (inventory <= g1) {
      *lower = mid;
      *upper = mid + 1;
      return;
    }
    if (fabs(inventory - g1) <= tol) {
      *lower = *upper = mid + 1;
      return;
    }
    if (mid + 1 == top) break;
    left = mid;
  }
  *err = LSMC_ERR_ARGUMENT;
  *lower = *upper = (inventory < grid[0]) ? 0 : n - 1;  // keep running on the nearest point; the flag fails the run
}

// The same result without the search when it cannot depend on the search path: on the fixed-spacing grid the bracketing
// interval is computed directly, and if the inventory is further than the tolerance from both of its end points the
// reference's bisection returns exactly that bracket whatever mid-points it visits.  Grid hits, tolerance cases and
// out-of-range inventories go through the reference's search.
//
// Within the tolerance of a grid point the reference's answer depends on which mid-points its search visits: for an
// inventory in (grid[j] - tol, grid[j]] it returns the point itself, (j, j), if mid = j is visited before mid = j - 1, and the
// bracket (j - 1, j) otherwise; above grid[j] within the tolerance it is always (j, j).  Which of the two mid-points comes
// first depends only on (n, j), so the host tabulates it per period by running the reference's search on grid[j] itself
// (first_hit[j] = 1: the point) and the kernel answers those cases without searching as well.  Paths sit on grid points all
// the time (an empty or full store is one), so the search itself — divergent, ~8 dependent iterations — was 23 % of the forward
// kernel's samples.  Anything else (outside the grid by more than the tolerance, one-point grids) still runs the search.
__device__ __forceinline__ void bracket_dev(const double *grid, const uint8_t *__restrict__ first_hit, int n, double inventory, double tol,
                                            double inv_spacing, int *lower, int *upper, int *err) {
  if (n >= 2) {
    int i = (int)((inventory - grid[0]) * inv_spacing);
    i = max(0, min(i, n - 2));
    if (inventory < grid[i] && i > 0)
      --i;
    else if (inventory > grid[i + 1] && i < n - 2)
      ++i;
    const double below = inventory - grid[i], above = grid[i + 1] - inventory;
    if (below > tol && above > tol) {
      *lower = i;
      *upper = i + 1;
      return;
    }
    if (below > 0.0 && below <= tol && above > tol) {  // just above grid[i]
      *lower = *upper = i;
      return;
    }
    if (below == 0.0 && above > tol) {  // on grid[i]
      const bool point = i == 0 || first_hit[i];
      *lower = point ? i : i - 1;
      *upper = i;
      return;
    }
    if (below > tol && above >= 0.0 && above <= tol) {  // on or just below grid[i + 1]
      const bool point = first_hit[i + 1];
      *lower = point ? i + 1 : i;
      *upper = i + 1;
      return;
    }
    if (below > tol && above < 0.0 && -above <= tol && i == n - 2) {  // just above the last grid point
      *lower = *upper = n - 1;
      return;
    }
  }
  bisect_dev(grid, n, inventory, tol, lower, upper, err);
}

__device__ __forceinline__ void cp_async8(void *dst_smem, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int KP>
__global__ void __launch_bounds__(kThreads, LSMC_FWD_MIN_CTAS) forward_kernel(ForwardArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int K = a.K, F = a.F;
  // dynamic smem: a.table_buffers x { DevPeriodParams | gamma [Gmax][K] | grid_next [Gmax] | rinv [K*K] dd | first_hit [first_pitch] }
  //               | warp staging | per-path state | powers
  const ForwardTableLayout lay = forward_table_layout(K, a.Gmax, a.first_pitch);
  unsigned char *s_tables = smem_raw;
  double *s_stage = reinterpret_cast<double *>(smem_raw + (size_t)a.table_buffers * lay.bytes);  // [kWarps][sums_width]
  double *s_inv = s_stage + kWarps * a.sums_width;                // [tiles_per_cta][kThreads] inventory carried by each path
  double *s_pv = s_inv + (size_t)a.tiles_per_cta * kThreads;       // [tiles_per_cta][kThreads] PV accumulated so far
  int32_t *s_pw = reinterpret_cast<int32_t *>(s_pv + (size_t)a.tiles_per_cta * kThreads);

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ MonomialProgram s_prog;
  for (int i = threadIdx.x; i < K * (1 + F); i += blockDim.x) s_pw[i] = a.powers[i];
  __syncthreads();
  if (threadIdx.x == 0) build_monomial_program(s_prog, s_pw, K, F);

  // The tables of a period (its parameters, the regression coefficients and the grid of the next period, R^-1, the first-visit
  // table of bracket_dev) are copied into shared memory asynchronously (cp.async), whole rows of the period-major device
  // arrays whatever the period's own grid size.  With two buffers the copy of period p + 1 is issued at the top of period p
  // and lands under its arithmetic, so a period costs one CTA barrier for the tables instead of three and no exposed load
  // latency: that is what a CTA with one or two path tiles (small path counts, many GPUs) spends its time on otherwise.
  auto stage_tables = [&](int p, int buf) {
    unsigned char *dst = s_tables + (size_t)buf * lay.bytes;
    const char *pp_src = reinterpret_cast<const char *>(a.periods + p);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPeriodParams) / 16); i += blockDim.x) cp_async16(dst + lay.params + 16 * i, pp_src + 16 * i);
    const double *gamma_src = a.gamma + (size_t)p * a.Gmax * K;
    for (int i = threadIdx.x; i < a.Gmax * K; i += blockDim.x) cp_async8(dst + lay.gamma + 8 * (size_t)i, gamma_src + i);
    const double *grid_src = a.grids + (size_t)(p + 1) * a.Gmax;
    for (int i = threadIdx.x; i < a.Gmax; i += blockDim.x) cp_async8(dst + lay.grid + 8 * (size_t)i, grid_src + i);
    const double *rinv_src = a.rinv + (size_t)p * K * K * 2;  // (hi, lo) pairs: the layout of dd
    for (int i = threadIdx.x; i < K * K; i += blockDim.x) cp_async16(dst + lay.rinv + 16 * (size_t)i, rinv_src + 2 * i);
    const uint8_t *first_src = a.bisect_first + (size_t)p * a.first_pitch;
    for (int i = threadIdx.x; i < a.first_pitch / 16; i += blockDim.x) cp_async16(dst + lay.first + 16 * (size_t)i, first_src + 16 * i);
    cp_async_commit();
  };

  // A CTA owns tiles blockIdx.x, blockIdx.x + gridDim.x, ... (at most a.tiles_per_cta of them).  Periods are the outer loop so
  // that the period's tables are staged once per CTA and period; the state a path carries from one period to the next
  // (inventory, PV so far) waits in shared memory.
  const int64_t num_tiles = (a.N + kThreads - 1) / kThreads;
  for (int t = 0; t < a.tiles_per_cta; ++t) {
    const int64_t n = (blockIdx.x + (int64_t)t * gridDim.x) * kThreads + threadIdx.x;
    const bool resume = a.p_begin > 0 && n < a.N;
    s_inv[t * kThreads + threadIdx.x] = resume ? a.final_inventory[n] : a.start_inventory;
    s_pv[t * kThreads + threadIdx.x] = resume ? a.pv_by_sim[n] : 0.0;
  }
  int err = 0;
  unsigned long long snapped = 0;  // inventories moved onto a limit (lsmc_results.snapped_inventories)

  const bool two = a.table_buffers > 1;
  if (two && a.p_begin < a.p_end) stage_tables(a.p_begin, 0);
  for (int p = a.p_begin; p < a.p_end; ++p) {
    const int buf = two ? (p - a.p_begin) & 1 : 0;
    if (!two) {
      __syncthreads();  // previous period's tables are no longer in use
      stage_tables(p, 0);
    }
    cp_async_wait_all();
    __syncthreads();  // this period's tables have landed; nobody reads the other buffer (period p - 1) or s_stage any more
    if (two && p + 1 < a.p_end) stage_tables(p + 1, buf ^ 1);
    const unsigned char *tab = s_tables + (size_t)buf * lay.bytes;
    const DevPeriodParams &pp = *reinterpret_cast<const DevPeriodParams *>(tab + lay.params);
    const double *s_gamma = reinterpret_cast<const double *>(tab + lay.gamma);
    const double *s_grid = reinterpret_cast<const double *>(tab + lay.grid);
    const dd *s_rinv = reinterpret_cast<const dd *>(tab + lay.rinv);
    const uint8_t *s_first = tab + lay.first;
    const int Gn = pp.grid_next_count;

    double sums[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // this thread's paths, tile after tile
    double qsum[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) qsum[k] = 0.0;

    BasisSpec b;
    b.mean_mode = pp.mean_mode;
    b.rinv = nullptr;
    b.spot = nullptr;
    if (!pp.mean_mode) {
      b.spot = a.spot + (size_t)pp.sim_row * a.sim_pitch;
#pragma unroll
      for (int i = 0; i < LSMC_MAX_FACTORS; ++i) b.factors[i] = (i < F) ? a.factors[i] + (size_t)pp.sim_row * a.sim_pitch : nullptr;
    }
    // the simulated inputs of the next tile are requested before the current tile's arithmetic
    PathInputs next;
    if (!pp.mean_mode && blockIdx.x < num_tiles) load_inputs(b, F, min((int64_t)blockIdx.x * kThreads + threadIdx.x, a.N - 1), next);

    for (int t = 0; t < a.tiles_per_cta; ++t) {
      const int64_t tile = blockIdx.x + (int64_t)t * gridDim.x;
      if (tile >= num_tiles) break;
      const int64_t n = tile * kThreads + threadIdx.x;
      const bool valid = n < a.N;
      double inventory = s_inv[t * kThreads + threadIdx.x];
      const PathInputs in = next;
      if (!pp.mean_mode && t + 1 < a.tiles_per_cta && tile + gridDim.x < num_tiles)
        load_inputs(b, F, min((tile + gridDim.x) * kThreads + threadIdx.x, a.N - 1), next);
      const double price = pp.mean_mode ? pp.forward_price : in.s;
      double q[KP];
#pragma unroll
      for (int k = 0; k < KP; ++k) q[k] = 0.0;
      if (valid) {
        if (pp.mean_mode) {
          q[0] = 1.0;
        } else {
          double X[KP];
          eval_monomials_from<KP>(in, s_prog, K, F, X);
          orthonormalise<KP>(X, s_rinv, K, q);
        }
      }

      // LSV.cs:445-508 for this path.  A breach of the inventory limits smaller than NumericalTolerance (rounding of
      // inv + v - loss after a constrained decision) is snapped onto the limit; the reference would throw (CS.cs:88-94).
      if (inventory < pp.min_inventory && pp.min_inventory - inventory <= a.tolerance) inventory = pp.min_inventory, snapped += valid;
      if (inventory > pp.max_inventory && inventory - pp.max_inventory <= a.tolerance) inventory = pp.max_inventory, snapped += valid;
      int perr = 0;  // this path's own status: threads beyond the last path run on placeholder state and must not report
      DevRange range = storage_range_dev(a, pp, inventory, &perr);
      const double loss = pp.loss * inventory;
      double decisions[LSMC_MAX_DECISIONS];
      const int D = decision_set_dev(range, inventory, loss, pp.next_space_lo, pp.next_space_hi, a.tolerance, a.extra_decisions,
                                     decisions, &perr);
      const double inv_cost_npv = (inventory * pp.inv_cost) * pp.inv_cost_df;
      double best = 0.0, best_consumed = 0.0, best_immediate = 0.0, best_volume = 0.0;
      int best_d = 0;
      for (int d = 0; d < D; ++d) {
        const double volume = decisions[d];
        const double after = inventory + volume - loss;
        const double consumed = (volume > 0.0 ? pp.cons_inj : pp.cons_wd) * fabs(volume);
        const double inject_withdraw_npv = (-volume * price) * pp.settle_df;
        const double consumed_npv = (-consumed * price) * pp.settle_df;
        const double cost_npv = volume > 0.0 ? (pp.inj_cost * volume) * pp.inj_df : (pp.wd_cost * fabs(-volume)) * pp.wd_df;
        const double immediate = ((inject_withdraw_npv - cost_npv) + consumed_npv) - inv_cost_npv;  // LSV.cs:473
        int lo, hi;
        bracket_dev(s_grid, s_first, Gn, after, a.tolerance, a.inv_grid_spacing, &lo, &hi, &perr);
        const double *gl = s_gamma + (size_t)lo * K;
        double c_lo = q[0] * gl[0];
#pragma unroll
        for (int k = 1; k < KP; ++k)
          if (k < K) c_lo = __fma_rn(q[k], gl[k], c_lo);
        double cont = c_lo;
        if (hi != lo) {  // InterpolateContinuationValue (LSV.cs:807-826)
          const double *gu = s_gamma + (size_t)hi * K;
          double c_hi = q[0] * gu[0];
#pragma unroll
          for (int k = 1; k < KP; ++k)
            if (k < K) c_hi = __fma_rn(q[k], gu[k], c_hi);
          const double lower_inv = s_grid[lo], upper_inv = s_grid[hi];
          const double lower_w = (upper_inv - after) / (upper_inv - lower_inv);
          const double upper_w = 1.0 - lower_w;
          cont = c_lo * lower_w + c_hi * upper_w;
        }
        const double total = immediate + cont;
        if (d == 0 || total > best) {
          best = total;
          best_d = d;
          best_consumed = consumed;
          best_immediate = immediate;
          best_volume = volume;
        }
      }
      const double next_inventory = inventory + best_volume - loss;  // LSV.cs:485
      if (valid && perr) err = perr;

      if (valid) {
        const size_t off = (size_t)(p - a.panel_row0) * a.N + n;
        if (a.p_inventory) a.p_inventory[off] = inventory;
        if (a.p_volume) a.p_volume[off] = best_volume;
        if (a.p_consumed) a.p_consumed[off] = best_consumed;
        if (a.p_loss) a.p_loss[off] = loss;
        if (a.p_net_volume) a.p_net_volume[off] = -best_volume - best_consumed;
        if (a.p_pv) a.p_pv[off] = best_immediate;
        if (a.p_decision) a.p_decision[off] = (double)best_d;
      }

      // per-period sums (LSV.cs:490-507, 510)
      if (valid) {
        sums[0] += -(best_volume + best_consumed) * price;
        sums[1] += best_volume;
        sums[2] += best_consumed;
        sums[3] += loss;
        sums[4] += best_immediate;
        sums[5] += inventory;
        s_pv[t * kThreads + threadIdx.x] += best_immediate;
      }
#pragma unroll
      for (int k = 0; k < KP; ++k) qsum[k] += q[k];
      s_inv[t * kThreads + threadIdx.x] = next_inventory;
    }

    // fixed-order reduction over the CTA
    int kidx;
    double total = warp_reduce8(sums, lane, kidx);
    if ((lane & 3) == 0) s_stage[warp * a.sums_width + kidx] = total;
    for (int half = 0; half < KP; half += 8) {
      double part[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) part[k] = (half + k < KP) ? qsum[(half + k < KP) ? half + k : 0] : 0.0;
      total = warp_reduce8(part, lane, kidx);
      if ((lane & 3) == 0 && half + kidx < K) s_stage[warp * a.sums_width + kForwardSums + half + kidx] = total;
    }
    __syncthreads();
    if (threadIdx.x < a.sums_width) {
      double t = s_stage[threadIdx.x];
      for (int w = 1; w < kWarps; ++w) t += s_stage[w * a.sums_width + threadIdx.x];
      a.step_sums[((size_t)p * gridDim.x + blockIdx.x) * a.sums_width + threadIdx.x] = t;
    }
  }
  for (int t = 0; t < a.tiles_per_cta; ++t) {
    const int64_t tile = blockIdx.x + (int64_t)t * gridDim.x;
    const int64_t n = tile * kThreads + threadIdx.x;
    if (tile < num_tiles && n < a.N) {
      const double inventory = s_inv[t * kThreads + threadIdx.x];
      a.pv_by_sim[n] = s_pv[t * kThreads + threadIdx.x];
      a.final_inventory[n] = inventory;
      if (a.p_inventory && a.p_end == a.P && a.panel_row0 == 0) a.p_inventory[(size_t)a.P * a.N + n] = inventory;
    }
  }
  if (err) atomicExch(a.error_flag, err);
  if (snapped) atomicAdd(a.snapped, snapped);  // integer sum: order-independent
}

// ------------------------------------------------------------------------------------------------------------------
// Deterministic sums
// ------------------------------------------------------------------------------------------------------------------
// out[block] = sum over this block's slice of (x[i] - shift)^power, power in {1,2}
__global__ void __launch_bounds__(256) sum_kernel(const double *__restrict__ x, int64_t n, double shift, int power, double *out) {
  __shared__ double s_w[8];
  const int64_t per_block = (n + gridDim.x - 1) / gridDim.x;
  const int64_t begin = (int64_t)blockIdx.x * per_block, end = min(begin + per_block, n);
  double acc = 0.0;
  for (int64_t i = begin + threadIdx.x; i < end; i += blockDim.x) {
    double d = x[i] - shift;
    acc += (power == 2) ? d * d : d;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(kFullMask, acc, off);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += s_w[w];
    out[blockIdx.x] = t;
  }
}

__global__ void combine_ranks_kernel(const double *__restrict__ all, int world, int count, int K, int mean_mode, double total_paths,
                                     double *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  double t = 0.0;
  for (int r = 0; r < world; ++r) t += all[(size_t)r * count + i];
  if (mean_mode) t = (i % K == 0) ? t / total_paths : 0.0;
  out[i] = t;
}

// out[p][j] = sum over CTAs of step_sums[p][cta][j], in CTA order
__global__ void reduce_step_sums_kernel(const double *__restrict__ in, int num_ctas, int width, double *out) {
  const int p = blockIdx.x;
  for (int j = threadIdx.x; j < width; j += blockDim.x) {
    double t = 0.0;
    for (int c = 0; c < num_ctas; ++c) t += in[((size_t)p * num_ctas + c) * width + j];
    out[(size_t)p * width + j] = t;
  }
}

template <int KP>
void launch_for_k(int which, const void *args, dim3 grid, dim3 block, size_t smem, cudaStream_t stream) {
  switch (which) {
    case 0:
      gram_kernel<KP><<<grid, block, 0, stream>>>(*static_cast<const GramArgs *>(args));
      break;
    case 1:
      cholesky_kernel<KP><<<grid, block, 0, stream>>>(*static_cast<const CholeskyArgs *>(args));
      break;
    case 2:
      if (static_cast<const BackwardStepArgs *>(args)->D == 3) {
        const BackwardStepArgs &ba = *static_cast<const BackwardStepArgs *>(args);
        if (ba.fused_payoff && !ba.terminal_mode) {
          ensure_dynamic_smem<backward_step_kernel<KP, 3, true>>(smem);
          backward_step_kernel<KP, 3, true><<<grid, block, smem, stream>>>(ba);
        } else {
          ensure_dynamic_smem<backward_step_kernel<KP, 3, false>>(smem);
          backward_step_kernel<KP, 3, false><<<grid, block, smem, stream>>>(ba);
        }
        break;
      }
      ensure_dynamic_smem<backward_step_kernel<KP, 0, false>>(smem);
      backward_step_kernel<KP, 0, false><<<grid, block, smem, stream>>>(*static_cast<const BackwardStepArgs *>(args));
      break;
    case 4:
      basis_kernel<KP><<<grid, block, 0, stream>>>(*static_cast<const BasisArgs *>(args));
      break;
    case 3:
      ensure_dynamic_smem<forward_kernel<KP>>(smem);
      forward_kernel<KP><<<grid, block, smem, stream>>>(*static_cast<const ForwardArgs *>(args));
      break;
  }
}

void dispatch(int KP, int which, const void *args, dim3 grid, dim3 block, size_t smem, cudaStream_t stream) {
  switch (KP) {
    case 1: launch_for_k<1>(which, args, grid, block, smem, stream); break;
    case 2: launch_for_k<2>(which, args, grid, block, smem, stream); break;
    case 3: launch_for_k<3>(which, args, grid, block, smem, stream); break;
    case 4: launch_for_k<4>(which, args, grid, block, smem, stream); break;
    case 5: launch_for_k<5>(which, args, grid, block, smem, stream); break;
    case 6: launch_for_k<6>(which, args, grid, block, smem, stream); break;
    case 7: launch_for_k<7>(which, args, grid, block, smem, stream); break;
    case 8: launch_for_k<8>(which, args, grid, block, smem, stream); break;
    case 12: launch_for_k<12>(which, args, grid, block, smem, stream); break;
    default: launch_for_k<16>(which, args, grid, block, smem, stream); break;
  }
}

}  // namespace

int padded_basis(int K) { return K <= 8 ? K : (K <= 12 ? 12 : 16); }

size_t backward_smem_bytes(int KP, int D) {
  const int acc = KP <= 8 ? 8 : 16, kc = (KP + 1) & ~1, tg = kMaxGridTile, pitch = 32 * kPathsPerThread + 4;
  return (size_t)tg * D * 64 + sizeof(double) * ((size_t)tg * D * kc + (size_t)kWarps * 8 * pitch + (size_t)kWarps * acc * pitch +
                                                (size_t)kWarps * acc * tg + tg) + sizeof(int32_t) * tg + 16;
}

size_t forward_smem_bytes(int K, int F, int Gmax, int first_pitch, int table_buffers, int sums_width, int tiles_per_cta) {
  return (size_t)table_buffers * forward_table_layout(K, Gmax, first_pitch).bytes + sizeof(double) * kWarps * sums_width +
         sizeof(double) * 2 * (size_t)tiles_per_cta * kThreads + sizeof(int32_t) * K * (1 + F) + 16;
}

void launch_gram(const GramArgs &a, int num_chunks, int num_slots, cudaStream_t stream) {
  dispatch(padded_basis(a.K), 0, &a, dim3(num_chunks, num_slots), dim3(kThreads), 0, stream);
}
void launch_cholesky(const CholeskyArgs &a, int num_slots, cudaStream_t stream) {
  dispatch(padded_basis(a.K), 1, &a, dim3(num_slots), dim3(64), 0, stream);
}
void launch_backward_step(const BackwardStepArgs &a, int gsplit, int64_t num_tiles, cudaStream_t stream) {
  const int KP = padded_basis(a.K);
  dispatch(KP, 2, &a, dim3(gsplit, (unsigned)num_tiles), dim3(kThreads), backward_smem_bytes(KP, a.D), stream);
}
void launch_basis(const BasisArgs &a, cudaStream_t stream) {
  static const int sm_count = [] {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
  }();
  const int64_t chunks = (a.pitch + 255) / 256;
  dispatch(padded_basis(a.K), 4, &a, dim3((unsigned)std::min<int64_t>(chunks, (int64_t)sm_count * 8)), dim3(256), 0, stream);
}
void launch_reduce_partials(const ReduceArgs &a, cudaStream_t stream) {
  const int threads = kReduceThreads;
  reduce_partials_kernel<<<a.G, threads, 0, stream>>>(a);
}
void launch_forward(const ForwardArgs &a, int num_ctas, cudaStream_t stream) {
  dispatch(padded_basis(a.K), 3, &a, dim3(num_ctas), dim3(kThreads),
           forward_smem_bytes(a.K, a.F, a.Gmax, a.first_pitch, a.table_buffers, a.sums_width, a.tiles_per_cta), stream);
}
void launch_combine_ranks(const double *all, int world, int count, int K, int mean_mode, double total_paths, double *out,
                          cudaStream_t stream) {
  combine_ranks_kernel<<<(count + 127) / 128, 128, 0, stream>>>(all, world, count, K, mean_mode, total_paths, out);
}
void launch_reduce_step_sums(const double *in, int P, int num_ctas, int width, double *out, cudaStream_t stream) {
  reduce_step_sums_kernel<<<P, 32, 0, stream>>>(in, num_ctas, width, out);
}
void launch_sum(const double *x, int64_t n, double shift, int power, double *out, int blocks, cudaStream_t stream) {
  sum_kernel<<<blocks, 256, 0, stream>>>(x, n, shift, power, out);
}

}  // namespace lsmc
