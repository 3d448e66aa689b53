// backward_common.cuh — pieces shared by the backward-induction kernels (kernels.cu: general kernel; backward_sweep.cu:
// three-decision sweep kernels).

#pragma once
#include <cstdint>

#include "device_types.hpp"

namespace lsmc {

struct __align__(16) DecisionA {  // one LDS.128 each
  double neg_volume, neg_consumed, cost_npv, lower_w;
};
struct __align__(16) DecisionR {  // upper weight and the offsets (in doubles) of the two rows of v_next
  double upper_w;
  uint32_t lower_off, upper_off;
};

constexpr int kWarpPaths = 32 * kPathsPerThread;  // 64 paths per warp
constexpr int kTilePitch = kWarpPaths + 4;        // row pitch of the warp-private tiles: conflict-free fragment loads

// d = a * b + c with separate accumulator-in and -out registers: the first mma of a chain reads its start value (a table
// constant held for the whole block) without a copy
__device__ __forceinline__ void dmma_m8n8k4_to(double &d0, double &d1, double a, double b, double c0, double c1) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
               : "=d"(d0), "=d"(d1)
               : "d"(a), "d"(b), "d"(c0), "d"(c1));
}
__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}


// Sim-independent coefficients of one decision: the next-grid interpolation folded into the regression coefficients
// (continuation value = q . c), written as a KC-long row.
template <int KC>
__device__ __forceinline__ void fold_coefficients(const DevDecision &e, const double *__restrict__ gamma, int K, double *__restrict__ c) {
  const double *gl = gamma + (size_t)e.lower * K, *gu = gamma + (size_t)e.upper * K;
#pragma unroll
  for (int k = 0; k < KC; ++k) c[k] = (k < K) ? (e.lower == e.upper ? gl[k] : gu[k] * e.upper_w + gl[k] * e.lower_w) : 0.0;
}


// One decision of one grid point for the two paths of a thread (LSV.cs:317-336): continuation value q . c, total NPV in
// the reference's association ((-v*S)*DF - cost + (-consumed*S)*DF) + continuation - inventoryCost, strict-> argmax
// (first maximum wins, SH.cs:206-221) and adj = -regressed + realised continuation (LSV.cs:330-332) where the realised
// value is the WeightedAverage of the v_next rows (LSV.cs:296-302, 828-836; an exact grid hit has weights (1,0) and
// reproduces the row bit for bit).
template <int KP>
__device__ __forceinline__ void evaluate_decision(bool first, const DecisionA &da, double upper_w, const double *__restrict__ c,
                                                  const double2 lo, const double2 up, const double (&q0)[KP], const double (&q1)[KP],
                                                  double price0, double price1, double df, double inv_cost, double &best0,
                                                  double &best1, double &adj0, double &adj1) {
  double cont0 = q0[0] * c[0], cont1 = q1[0] * c[0];
#pragma unroll
  for (int k = 1; k < KP; ++k) {
    cont0 = __fma_rn(q0[k], c[k], cont0);
    cont1 = __fma_rn(q1[k], c[k], cont1);
  }
  const double t0 = ((((da.neg_volume * price0) * df - da.cost_npv) + (da.neg_consumed * price0) * df) + cont0) - inv_cost;
  const double t1 = ((((da.neg_volume * price1) * df - da.cost_npv) + (da.neg_consumed * price1) * df) + cont1) - inv_cost;
  const double a0 = -cont0 + (up.x * upper_w + lo.x * da.lower_w);
  const double a1 = -cont1 + (up.y * upper_w + lo.y * da.lower_w);
  if (first || t0 > best0) {
    best0 = t0;
    adj0 = a0;
  }
  if (first || t1 > best1) {
    best1 = t1;
    adj1 = a1;
  }
}

// Rows of v_next one grid point needs: (lower, upper) per decision, both paths of the thread in one 16-byte load each.
struct Rows {
  double2 lo[3], up[3];
};

__device__ __forceinline__ void load_rows(const DecisionR *__restrict__ pb, int nd, const double *__restrict__ vbase, Rows &r) {
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (d < nd) {
      r.lo[d] = *reinterpret_cast<const double2 *>(vbase + pb[d].lower_off);
      r.up[d] = *reinterpret_cast<const double2 *>(vbase + pb[d].upper_off);
    }
  }
}

// One grid point with at most 3 decisions for the thread's two paths, its v_next rows already in registers.
template <int KP, int KC>
__device__ __forceinline__ void point(const DecisionA *__restrict__ pa, const DecisionR *__restrict__ pb, const double *__restrict__ pc,
                                      int nd, double inv_cost, const Rows &r, const double (&q0)[KP], const double (&q1)[KP],
                                      double price0, double price1, double df, double &v0, double &v1) {
  double best0 = 0.0, best1 = 0.0, adj0 = 0.0, adj1 = 0.0;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (d < nd) {
      const DecisionA da = pa[d];
      evaluate_decision<KP>(d == 0, da, pb[d].upper_w, pc + d * KC, r.lo[d], r.up[d], q0, q1, price0, price1, df, inv_cost, best0,
                            best1, adj0, adj1);
    }
  }
  v0 = best0 + adj0;  // LSV.cs:333
  v1 = best1 + adj1;
}

// Fused-payoff form of a decision.  The reference's five roundings of the immediate cash flow,
// ((-v*S)*DF - cost) + (-consumed*S)*DF ... - inventoryCost, collapse into one FMA S*price_coef + offset with
// price_coef = (-v - consumed)*DF and offset = -(cost + inventoryCost) folded per decision when the tables are staged, and
// the value carried for the winning decision is payoff + realised continuation directly instead of
// (payoff + regressed) + (realised - regressed).  Same real-number result; differs from the reference's roundings by a
// few ulp of the payoff, far below the 1e-9 parity tolerance.  13 FP64 operations per decision and path instead of 20.
struct __align__(16) DecisionF {
  double price_coef, offset, lower_w, upper_w;
};
static_assert(sizeof(DecisionF) == sizeof(DecisionA), "the fused table re-uses the DecisionA slots");

template <int KP, int KC>
__device__ __forceinline__ void point_fused(const DecisionF *__restrict__ pf, const double *__restrict__ pc, int nd, const Rows &r,
                                            const double (&q0)[KP], const double (&q1)[KP], double price0, double price1, double &v0,
                                            double &v1) {
  double best0 = 0.0, best1 = 0.0;
  v0 = v1 = 0.0;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (d < nd) {
      const DecisionF f = pf[d];
      const double *c = pc + d * KC;
      double cont0 = q0[0] * c[0], cont1 = q1[0] * c[0];
#pragma unroll
      for (int k = 1; k < KP; ++k) {
        cont0 = __fma_rn(q0[k], c[k], cont0);
        cont1 = __fma_rn(q1[k], c[k], cont1);
      }
      const double pay0 = __fma_rn(price0, f.price_coef, f.offset), pay1 = __fma_rn(price1, f.price_coef, f.offset);
      const double t0 = pay0 + cont0, t1 = pay1 + cont1;  // what the argmax is taken over (LSV.cs:317-327)
      const double w0 = pay0 + __fma_rn(r.up[d].x, f.upper_w, r.lo[d].x * f.lower_w);
      const double w1 = pay1 + __fma_rn(r.up[d].y, f.upper_w, r.lo[d].y * f.lower_w);
      if (d == 0 || t0 > best0) {
        best0 = t0;
        v0 = w0;
      }
      if (d == 0 || t1 > best1) {
        best1 = t1;
        v1 = w1;
      }
    }
  }
}

// Any number of decisions (ExtraDecisions > 0): rows are loaded as they are needed.
template <int KP, int KC>
__device__ __forceinline__ void point_generic(const DecisionA *__restrict__ pa, const DecisionR *__restrict__ pb,
                                              const double *__restrict__ pc, int nd, double inv_cost, const double *__restrict__ vbase,
                                              const double (&q0)[KP], const double (&q1)[KP], double price0, double price1, double df,
                                              double &v0, double &v1) {
  double best0 = 0.0, best1 = 0.0, adj0 = 0.0, adj1 = 0.0;
  for (int d = 0; d < nd; ++d) {
    const DecisionA da = pa[d];
    const DecisionR db = pb[d];
    const double2 lo = *reinterpret_cast<const double2 *>(vbase + db.lower_off);
    const double2 up = *reinterpret_cast<const double2 *>(vbase + db.upper_off);
    evaluate_decision<KP>(d == 0, da, db.upper_w, pc + (size_t)d * KC, lo, up, q0, q1, price0, price1, df, inv_cost, best0, best1, adj0,
                          adj1);
  }
  v0 = best0 + adj0;
  v1 = best1 + adj1;
}

}  // namespace lsmc
