// backward_sweep.cu — the backward-induction step (LSV.cs:240-338) for the usual three-decision table
// (inject / hold / withdraw, no extra decisions).  kernels.cu keeps the general kernel (any table width, terminal step).
//
// Work layout.  A CTA owns 128 paths and a contiguous range of inventory-grid points.  Its 8 warps are 2 path groups x 4
// grid slots: every lane owns two adjacent paths of its group (double2 loads), and the four warps of a group take the grid
// points g = slot (mod 4).  The rows of v_next a decision interpolates between slide upwards with g, so four warps working
// on neighbouring grid points of the same paths share one sliding window of rows: the window of a group is a few KB and
// stays in L1, where the same 16 warps sweeping their own path tiles would each keep their own window and spill to L2.
//
// All sim-independent operands of a grid point — decision payoffs, interpolation weights and row offsets, and the
// regression coefficients with the next-grid interpolation folded in — are prepared once per step by fold_tables_kernel
// into one record per grid point; the sweep kernel copies records into a shared-memory ring with cp.async, two sub-tiles
// ahead, so no warp waits for them.
//
// The regression sums of the next backward step, sum_n q_prev[k][n] * V[g][n] (the right-hand side of LSV.cs:199-212 in
// the orthonormal basis), are a skinny FP64 GEMM run on the tensor path (mma.sync m8n8k4 f64) once per block of 8 grid
// points: the warps publish their V values to a double-buffered shared tile, meet at one barrier, and each multiplies its
// 16-path slice, with its slice of q_prev held in registers.  Slices are combined in a fixed order: deterministic.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "backward_common.cuh"
#include "device_types.hpp"
#include "kernels.hpp"
#include "reduce_partials.cuh"
#include "mma_records.cuh"
#include "sweep_common.cuh"

#ifndef LSMC_SWEEP_PF
#define LSMC_SWEEP_PF 8  // rows ahead of the window that are requested into L2
#endif
#ifndef LSMC_SWEEP_MIN_CTAS
#define LSMC_SWEEP_MIN_CTAS 2
#endif

namespace lsmc {

namespace {

constexpr int kSweepThreads = 256;
constexpr int kSweepWarps = kSweepThreads / 32;
constexpr int kSlots = 4;                             // warps sharing a path group, one grid point each (mod 4)
constexpr int kGroupPaths = 64;                       // 32 lanes x double2
constexpr int kBlockG = 8;                            // grid points per block = n of the m8n8k4 tile
constexpr int kTabSub = 32;                           // records per staged sub-tile
constexpr int kTabRing = 2 * kTabSub;                 // records resident in shared memory
constexpr int kVPitch = kSweepTilePaths + 4;          // conflict-free fragment loads (pitch = 8 banks mod 32)
static_assert(kSweepTilePaths == (kSweepWarps / kSlots) * kGroupPaths, "two path groups per CTA");

// Everything one grid point needs that does not depend on the path.  Always three decisions: a grid point with fewer
// carries copies of its first decision in the unused slots, which can never win the strict > comparison (SH.cs:206-221),
// so the sweep has no per-decision branches.
template <int KC>
struct __align__(16) PointTable {
  DecisionA a[3];     // fused payoff: {price_coef, offset, lower_w, upper_w} (DecisionF); reference order: DecisionA
  double c[3][KC];    // regression coefficients with the next-grid interpolation folded in
  double upper_w[3];  // reference order only
  double inv_cost;    // reference order only
  uint32_t lower_row[3], upper_row[3];  // rows of v_next the decision interpolates between
  uint32_t ahead_row;                   // row to request into L2 now: a few rows above this point's window
  int32_t nd;
};

__device__ __forceinline__ void cp_async_16(void *dst_smem, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int KC>
__device__ __forceinline__ void fold_item(const FoldArgs &f, int i) {  // i = 3 g + d
  const int g = i / 3, d = i - 3 * g;
  PointTable<KC> *t = reinterpret_cast<PointTable<KC> *>(f.out) + g;
  const int nd = f.num_decisions[g];
  const double inv = f.inventory_cost[g];
  if (d == 0) {
    t->inv_cost = inv;
    t->nd = nd;
    const uint32_t top = (uint32_t)f.decisions[(size_t)g * 3 + (nd - 1)].upper + (uint32_t)LSMC_SWEEP_PF;
    t->ahead_row = f.Gn > 0 ? min(top, (uint32_t)(f.Gn - 1)) : 0u;
  }
  const DevDecision e = f.decisions[(size_t)g * 3 + (d < nd ? d : 0)];
  if (f.fused_payoff)
    t->a[d] = DecisionA{(-e.volume - e.consumed) * f.settle_df, -(e.cost_npv + inv), e.lower_w, e.upper_w};
  else
    t->a[d] = DecisionA{-e.volume, -e.consumed, e.cost_npv, e.lower_w};
  t->upper_w[d] = e.upper_w;
  // direct-load kernel: row indices; row-ring kernel: byte offsets of the rows' slots in the shared-memory ring
  t->lower_row[d] = f.ring_rows > 0 ? (uint32_t)(e.lower % f.ring_rows) * (uint32_t)(kSweepTilePaths * 8) : (uint32_t)e.lower;
  t->upper_row[d] = f.ring_rows > 0 ? (uint32_t)(e.upper % f.ring_rows) * (uint32_t)(kSweepTilePaths * 8) : (uint32_t)e.upper;
  // gamma may have been written by other CTAs of the same launch (reduce_fold_kernel): read it through L2
  const double *gl = f.gamma + (size_t)e.lower * f.K, *gu = f.gamma + (size_t)e.upper * f.K;
#pragma unroll
  for (int k = 0; k < KC; ++k) {
    double c = 0.0;
    if (k < f.K) c = (e.lower == e.upper) ? __ldcg(gl + k) : __ldcg(gu + k) * e.upper_w + __ldcg(gl + k) * e.lower_w;
    t->c[d][k] = c;
  }
}

template <int KC>
__global__ void fold_tables_kernel(FoldArgs f) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= fold_item_count(f)) return;
  if (f.format == 1)
    fold_item_mma(f, i);
  else
    fold_item<KC>(f, i);
}

// The per-step reduction and the table fold in one launch: every CTA reduces one next-grid point; the CTA that finishes
// last (device-wide counter) has all of gamma in front of it and folds the point tables of the step.  Saves a dependent
// launch that would otherwise queue behind the basis kernel of the second stream.
template <int KC>
__global__ void __launch_bounds__(kReduceThreads) reduce_fold_kernel(ReduceArgs r, FoldArgs f, unsigned int *counter) {
  asm volatile("griddepcontrol.launch_dependents;");  // the sweep kernel behind this one may start its prologue now (it waits for the records)
  reduce_partials_body(r);
  __shared__ bool s_last;
  __threadfence();  // this CTA's gamma row is visible device-wide before it is counted
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int items = fold_item_count(f);
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    if (f.format == 1)
      fold_item_mma(f, i);
    else
      fold_item<KC>(f, i);
  }
  if (threadIdx.x == 0) *counter = 0u;  // ready for the next launch
}

// ------------------------------------------------------------------------------------------------------------------
// Sharded runs: the per-step chain  reduce partials -> exchange between ranks -> sum in rank order -> fold tables  as ONE
// launch.  Every CTA reduces one next-grid point and stores its K sums straight into every rank's inbox (peer memory over
// NVLink; this rank's own inbox included).  The CTA that finishes last publishes this rank's arrival flag on every peer, waits
// until every rank's flag has arrived here, adds the world x G' x K sums in rank order (bit-identical on every rank) into
// gamma and folds the step's tables.  Only that one CTA ever waits, after this rank's own data and flags are out, so ranks
// cannot dead-lock on one another; a peer that never arrives (it failed or was cancelled) is given up on after timeout_ns
// and reported through the error flag instead of hanging the GPU.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <int KC>
__global__ void __launch_bounds__(kReduceThreads) exchange_fold_kernel(ReduceArgs r, FoldArgs f, XchgArgs x, unsigned int *counter) {
  asm volatile("griddepcontrol.launch_dependents;");  // as in reduce_fold_kernel
  const int count = r.G * r.K;
  const size_t data_off = xchg_data_offset(x.world);
  const size_t slot = ((size_t)(x.seq & 1) * x.world + x.rank) * x.capacity;  // this rank's slot in every inbox
  __shared__ double s_sums[LSMC_MAX_BASIS];
  {  // local sums of this CTA's grid point (mean mode is applied after the ranks are combined)
    ReduceArgs local = r;
    local.mean_mode = 0;
    local.out = nullptr;
    reduce_partials_body(local, s_sums);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < r.K * x.world; i += blockDim.x) {
    const int peer = i / r.K, k = i - peer * r.K;
    reinterpret_cast<double *>(x.peers[peer] + data_off)[slot + (size_t)blockIdx.x * r.K + k] = s_sums[k];
  }
  __shared__ bool s_last;
  __threadfence_system();  // this CTA's sums are visible on every peer before it is counted
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence_system();
  if (threadIdx.x < x.world)
    st_release_sys(reinterpret_cast<unsigned long long *>(x.peers[threadIdx.x] + (size_t)x.rank * kXchgFlagStride), x.seq);
  if (threadIdx.x < x.world) {
    const unsigned long long *flag = reinterpret_cast<const unsigned long long *>(x.peers[x.rank] + (size_t)threadIdx.x * kXchgFlagStride);
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys(flag) < x.seq) {
      if (*reinterpret_cast<volatile int32_t *>(x.error_flag) != 0) break;  // an earlier step already gave up: do not wait again
      if (global_timer_ns() - t0 > x.timeout_ns) {
        atomicCAS(x.error_flag, 0, (int)LSMC_ERR_CUDA);
        break;
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
  __threadfence_system();
  // gamma = sum over ranks, in rank order; the inbox was written by other GPUs: read it past L1
  const double *inbox = reinterpret_cast<const double *>(x.peers[x.rank] + data_off) + (size_t)(x.seq & 1) * x.world * x.capacity;
  for (int i = threadIdx.x; i < count; i += blockDim.x) {
    double t = 0.0;
    for (int s = 0; s < x.world; ++s) t += __ldcv(inbox + (size_t)s * x.capacity + i);
    if (r.mean_mode) t = (i % r.K == 0) ? t / (double)r.total_paths : 0.0;
    r.out[i] = t;
  }
  __threadfence();
  __syncthreads();
  const int items = f.G > 0 ? fold_item_count(f) : 0;
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    if (f.format == 1)
      fold_item_mma(f, i);
    else
      fold_item<KC>(f, i);
  }
  if (threadIdx.x == 0) *counter = 0u;  // ready for the next launch
}

// 16-byte load of this thread's two columns of row `row` (v_next is read-only for the whole launch)
__device__ __forceinline__ double2 load_row(const char *base, uint32_t row, uint32_t pitch_bytes) {
  double2 v;
  asm volatile("{\n\t.reg .u64 p;\n\tmad.wide.u32 p, %2, %3, %4;\n\tld.global.v2.f64 {%0, %1}, [p];\n\t}"
      : "=d"(v.x), "=d"(v.y)
      : "r"(row), "r"(pitch_bytes), "l"(base));
  return v;
}

template <int KC>
__device__ __forceinline__ void load_rows3(const PointTable<KC> *__restrict__ t, const char *__restrict__ vbase, uint32_t pitch_bytes,
                                           Rows &r) {
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    r.lo[d] = load_row(vbase, t->lower_row[d], pitch_bytes);
    r.up[d] = load_row(vbase, t->upper_row[d], pitch_bytes);
  }
}

// One grid point for the thread's two paths (LSV.cs:310-336).  `r` holds the rows of v_next this point interpolates between;
// as soon as a decision has consumed its pair of rows the same registers are re-loaded with the rows that decision needs at
// the warp's next grid point (record tn), so every load has most of a grid point's arithmetic to complete behind.
template <int KP, int KC, bool FUSED>
__device__ __forceinline__ void evaluate3(const PointTable<KC> *__restrict__ t, const PointTable<KC> *__restrict__ tn,
                                          const char *__restrict__ vbase, uint32_t pitch_bytes, Rows &r, const double (&q0)[KP],
                                          const double (&q1)[KP], double price0, double price1, double df, double &v0, double &v1) {
  double best0 = 0.0, best1 = 0.0, adj0 = 0.0, adj1 = 0.0;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    if (FUSED) {
      const DecisionF f = *reinterpret_cast<const DecisionF *>(&t->a[d]);
      const double *c = t->c[d];
      const double real0 = __fma_rn(r.up[d].x, f.upper_w, r.lo[d].x * f.lower_w);  // realised continuation value (LSV.cs:296-302)
      const double real1 = __fma_rn(r.up[d].y, f.upper_w, r.lo[d].y * f.lower_w);
      r.lo[d] = load_row(vbase, tn->lower_row[d], pitch_bytes);
      r.up[d] = load_row(vbase, tn->upper_row[d], pitch_bytes);
      double cont0 = q0[0] * c[0], cont1 = q1[0] * c[0];
#pragma unroll
      for (int k = 1; k < KP; ++k) {
        cont0 = __fma_rn(q0[k], c[k], cont0);
        cont1 = __fma_rn(q1[k], c[k], cont1);
      }
      const double pay0 = __fma_rn(price0, f.price_coef, f.offset), pay1 = __fma_rn(price1, f.price_coef, f.offset);
      const double t0 = pay0 + cont0, t1 = pay1 + cont1;  // what the argmax is taken over (LSV.cs:317-327)
      const double w0 = pay0 + real0, w1 = pay1 + real1;
      if (d == 0 || t0 > best0) {
        best0 = t0;
        v0 = w0;
      }
      if (d == 0 || t1 > best1) {
        best1 = t1;
        v1 = w1;
      }
    } else {
      evaluate_decision<KP>(d == 0, t->a[d], t->upper_w[d], t->c[d], r.lo[d], r.up[d], q0, q1, price0, price1, df, t->inv_cost, best0,
                            best1, adj0, adj1);
      r.lo[d] = load_row(vbase, tn->lower_row[d], pitch_bytes);
      r.up[d] = load_row(vbase, tn->upper_row[d], pitch_bytes);
    }
  }
  if (!FUSED) {
    v0 = best0 + adj0;  // LSV.cs:333
    v1 = best1 + adj1;
  }
}

template <int KP, bool FUSED>
__global__ void __launch_bounds__(kSweepThreads, LSMC_SWEEP_MIN_CTAS) backward_sweep_kernel(BackwardStepArgs a) {
  constexpr int KC = (KP + 1) & ~1;
  constexpr int MB = KP <= 8 ? 1 : 2;  // 8-row blocks of the regression sums
  constexpr int ACC = 8 * MB;          // width of a stored partial-sum row
  using Table = PointTable<KC>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Table *s_tab = reinterpret_cast<Table *>(smem_raw);              // [kTabRing]
  double *s_v = reinterpret_cast<double *>(s_tab + kTabRing);      // [2][kBlockG][kVPitch]
  double *s_acc = s_v + 2 * kBlockG * kVPitch;                     // [2][kSweepWarps][kBlockG][ACC]

  const int K = a.K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = warp & (kSlots - 1), group = warp / kSlots;
  const int64_t tile = blockIdx.y;
  const int64_t n0 = tile * kSweepTilePaths + group * kGroupPaths + 2 * lane;
  const bool valid0 = n0 < a.N;
  // Threads beyond the last path work on a clamped (valid, even) column pair and never store; the odd column of a
  // half-valid pair is the padding column of the row buffers, which holds finite values throughout (zero basis rows,
  // zero terminal values).  Neither reaches the regression sums: the q_prev fragments below are zero beyond N.
  const int64_t c0 = valid0 ? n0 : ((a.N - 1) & ~int64_t(1));
  const int64_t c1 = min(c0 + 1, a.N - 1);

  // this CTA's grid points
  const int total_blocks = (a.G + kBlockG - 1) / kBlockG;
  const int bbeg = blockIdx.x * a.subtiles_per_cta;
  const int nblocks = min(bbeg + a.subtiles_per_cta, total_blocks) - bbeg;
  if (nblocks <= 0) return;
  const int gbeg = bbeg * kBlockG;
  const int glimit = min(a.G, (bbeg + nblocks) * kBlockG) - gbeg;  // grid points of this CTA, addressed 0..glimit-1 below

  const Table *tab = reinterpret_cast<const Table *>(a.point_tables) + gbeg;
  auto stage = [&](int sub) {  // records [32 sub, 32 sub + 32) of this CTA's range into their half of the ring
    const int first = sub * kTabSub;
    const int count = min(kTabSub, glimit - first);
    if (count <= 0) return;
    const char *src = reinterpret_cast<const char *>(tab + first);
    char *dst = reinterpret_cast<char *>(s_tab + (sub & 1) * kTabSub);
    const int chunks = count * (int)(sizeof(Table) / 16);
    for (int i = threadIdx.x; i < chunks; i += kSweepThreads) cp_async_16(dst + 16 * i, src + 16 * i);
  };
  stage(0);
  stage(1);

  double price0, price1;
  if (a.spot) {
    price0 = a.spot[c0];
    price1 = a.spot[c1];
  } else {
    price0 = price1 = a.const_price;
  }
  // orthonormal basis row of this period for the thread's two paths (continuation value = q . c)
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
  if (!a.q_this) q0[0] = q1[0] = 1.0;  // mean mode: continuation value = gamma[g][0]

  // A operand of the regression GEMM for this warp's 16-path slice of the CTA: q_prev[k][slice], fragment layout of
  // mma.m8n8k4 (row = lane / 4 -> k, col = lane % 4 -> path inside a chunk of 4).  Paths beyond N contribute zero.
  const int row = lane >> 2, col = lane & 3;
  double afrag[MB][4];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = mb * 8 + row;
      const int64_t p = tile * kSweepTilePaths + 16 * warp + 4 * j + col;
      double v = 0.0;
      if (a.accumulate && k < K && p < a.N) v = a.q_prev ? a.q_prev[(size_t)k * a.pitch + p] : (k == 0 ? 1.0 : 0.0);
      afrag[mb][j] = v;
    }
  }

  const double df = a.settle_df;
  const uint32_t pitch_bytes = (uint32_t)a.pitch * 8u;
  // this thread's two columns of every v_next row; produced by a wide multiply-add so that it lives in an aligned register
  // pair and row addresses are a single IMAD.WIDE each
  const char *vbase = row_address(reinterpret_cast<const char *>(a.v_next), (uint32_t)c0, 8u);
  // this warp's grid points are slot, slot + 4, ...: two per block.  Points past the end of the CTA's range are evaluated
  // on the last record instead (and not stored), so the loop body is branch-free.
  const int last = glimit - 1;
  char *vout = reinterpret_cast<char *>(a.v_this + (size_t)gbeg * a.pitch + n0) + (uint64_t)slot * pitch_bytes;
  double *my_v = s_v + (size_t)slot * kVPitch + group * kGroupPaths + 2 * lane;   // + buffer; second point kSlots rows further
  const double *my_b = s_v + (size_t)row * kVPitch + 16 * warp + col;             // B fragments of this warp's slice
  double *my_c = s_acc + (size_t)warp * kBlockG * ACC + (2 * col) * ACC + row;    // C fragment destination

  auto combine = [&](int b) {  // partial sums of block b: the 8 warps' slices added in warp order
    if (threadIdx.x < kBlockG * ACC) {
      const int g8 = threadIdx.x / ACC, k = threadIdx.x - g8 * ACC;
      const int g = b * kBlockG + g8;
      if (g < glimit) {
        const double *src = s_acc + (size_t)(b & 1) * kSweepWarps * kBlockG * ACC + g8 * ACC + k;
        double sum = src[0];
#pragma unroll
        for (int w = 1; w < kSweepWarps; ++w) sum += src[(size_t)w * kBlockG * ACC];
        a.partials[((size_t)(gbeg + g) * gridDim.y + tile) * ACC + k] = sum;
      }
    }
  };

  cp_async_wait_all();
  __syncthreads();

  Rows rows;
  load_rows3(s_tab + min(slot, last), vbase, pitch_bytes, rows);
  for (int b = 0; b < nblocks; ++b) {
    // this warp's two grid points of the block, and its first one of the next block (rows requested one point ahead)
    const int gA = b * kBlockG + slot, gB = gA + kSlots, gA2 = gA + kBlockG;
    const Table *tA = s_tab + (min(gA, last) & (kTabRing - 1));
    const Table *tB = s_tab + (min(gB, last) & (kTabRing - 1));
    const Table *tA2 = s_tab + (min(gA2, last) & (kTabRing - 1));
    double vA0, vA1, vB0, vB1;
#if LSMC_SWEEP_PF > 0
    // the window's top row climbs about one row per grid point: ask L2 for the rows it reaches a few points from now
    asm volatile("prefetch.global.L2 [%0];" ::"l"(row_address(vbase, tA->ahead_row, pitch_bytes)));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(row_address(vbase, tB->ahead_row, pitch_bytes)));
#endif
    evaluate3<KP, KC, FUSED>(tA, tB, vbase, pitch_bytes, rows, q0, q1, price0, price1, df, vA0, vA1);
    evaluate3<KP, KC, FUSED>(tB, tA2, vbase, pitch_bytes, rows, q0, q1, price0, price1, df, vB0, vB1);
    if (valid0) {
      if (gA <= last) *reinterpret_cast<double2 *>(vout) = make_double2(vA0, vA1);
      if (gB <= last) *reinterpret_cast<double2 *>(vout + (uint64_t)kSlots * pitch_bytes) = make_double2(vB0, vB1);
    }
    vout += (uint64_t)kBlockG * pitch_bytes;
    const int buf = b & 1;
    if (a.accumulate) {
      double *vt = my_v + buf * (kBlockG * kVPitch);
      *reinterpret_cast<double2 *>(vt) = make_double2(vA0, vA1);
      *reinterpret_cast<double2 *>(vt + kSlots * kVPitch) = make_double2(vB0, vB1);
    }
    // records of the next sub-tile must be visible one block before its first block, whose rows are requested in the
    // last block of this sub-tile
    if ((b & 3) == 2) cp_async_wait_all();
    __syncthreads();
    if ((b & 3) == 3) stage((b >> 2) + 2);  // every warp is done with the sub-tile that ends here: refill its half
    if (a.accumulate) {
      // C[k][g] = sum over the warp's 16 paths of q_prev[k][n] * V[g][n]; C fragment: row -> k, columns 2 col + {0,1} -> g
      const double *vb = my_b + buf * (kBlockG * kVPitch);
      double *dst = my_c + buf * (kSweepWarps * kBlockG * ACC);
#pragma unroll
      for (int mb = 0; mb < MB; ++mb) {
        double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma_m8n8k4(acc0, acc1, afrag[mb][j], vb[4 * j]);
        dst[mb * 8] = acc0;
        dst[ACC + mb * 8] = acc1;
      }
      if (b > 0) combine(b - 1);  // complete since the barrier above
    }
  }
  if (a.accumulate) {
    __syncthreads();
    combine(nblocks - 1);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Row-ring variant.  The profile of the kernel above shows the L1/shared-memory data pipe as the limiter: every warp reads
// each grid point's record with broadcast loads (8 useful bytes per wavefront) and its rows of v_next through the same
// pipe.  This variant halves the record traffic per path and removes the global loads from the evaluating threads:
//   * a CTA is ONE group of 128 paths x 4 grid slots (4 warps); every lane owns FOUR paths (two adjacent pairs 64 columns
//     apart, so every 16-byte access of a warp is one contiguous 512-byte run), and one set of record loads serves 128 paths;
//   * the window of v_next rows that a block of 8 grid points reads ([lo, hi], SweepBlock) lives in a shared-memory ring of
//     R rows x 128 paths.  While block b is evaluated, one warp issues "generation" b+1: a bulk asynchronous copy (TMA 1-D,
//     cp.async.bulk) per 1 KB row that block b+1 adds, plus, every other block, one bulk copy of the next 16 point records;
//     they complete by byte count on one of two alternating mbarriers, which everybody waits on before starting the block.
//     There is no CTA barrier in the loop: the hand-offs between warps (block evaluated, GEMM slices written) are mbarriers
//     that producers arrive on and only consumers wait on.  No thread holds registers for data in flight and the copies do
//     not cross the LSU data pipe.  Every row comes from HBM/L2 once per CTA and is read back with conflict-free
//     LDS.128 at offsets the fold kernel precomputed.  (An extra L2 look-ahead was measured and is slower: off.)
// The host uses it when the windows slide upwards and fit (lsmc_api.cu), otherwise the direct-load kernel above.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kRingThreads = 128;
constexpr int kRingWarps = kRingThreads / 32;          // = kSlots: one warp per grid slot
constexpr int kLanePaths = 4;                          // paths per lane: two adjacent pairs ...
constexpr int kPairGap = kSweepTilePaths / 2;          // ... this many columns apart
constexpr int kRowBytes = kSweepTilePaths * 8;
constexpr int kRingTabSub = 16, kRingTabRing = 2 * kRingTabSub;
constexpr int kVBufs = 4;    // regression value tiles in flight (the warps of a CTA may be a block apart, the GEMM runs a block behind)
constexpr int kAccBufs = 3;  // regression-sum slice sets in flight
static_assert(kRingWarps == kSlots && 32 * kLanePaths == kSweepTilePaths, "one 128-path group per CTA");
#ifndef LSMC_RING_PF_AHEAD
#define LSMC_RING_PF_AHEAD 0  // generations ahead whose rows are requested into L2 (measured: 0 is fastest, the bulk copies cover the latency)
#endif
#ifndef LSMC_RING_MIN_CTAS
#define LSMC_RING_MIN_CTAS 3
#endif

// One grid point for the lane's four paths; rows from the shared-memory ring (LSV.cs:310-336).
template <int KP, int KC, bool FUSED>
__device__ __forceinline__ void evaluate4_ring(const PointTable<KC> *__restrict__ t, const char *__restrict__ my_rows,
                                               const double (&q)[kLanePaths][KP], const double (&price)[kLanePaths], double df,
                                               double (&v)[kLanePaths]) {
  if (FUSED) {
    double best[kLanePaths];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const DecisionF f = *reinterpret_cast<const DecisionF *>(&t->a[d]);
      const double *c = t->c[d];
      const double2 lo01 = *reinterpret_cast<const double2 *>(my_rows + t->lower_row[d]);
      const double2 lo23 = *reinterpret_cast<const double2 *>(my_rows + t->lower_row[d] + kPairGap * 8);
      // realised continuation value (LSV.cs:296-302).  A decision that lands exactly on a grid point of the next period has
      // lower == upper and weights (1, 0): the interpolation returns the row itself, bit for bit, so it is not re-read.
      double real[kLanePaths] = {lo01.x, lo01.y, lo23.x, lo23.y};
      if (t->lower_row[d] != t->upper_row[d]) {  // warp-uniform
        const double2 up01 = *reinterpret_cast<const double2 *>(my_rows + t->upper_row[d]);
        const double2 up23 = *reinterpret_cast<const double2 *>(my_rows + t->upper_row[d] + kPairGap * 8);
        real[0] = __fma_rn(up01.x, f.upper_w, real[0] * f.lower_w);
        real[1] = __fma_rn(up01.y, f.upper_w, real[1] * f.lower_w);
        real[2] = __fma_rn(up23.x, f.upper_w, real[2] * f.lower_w);
        real[3] = __fma_rn(up23.y, f.upper_w, real[3] * f.lower_w);
      }
      double cont[kLanePaths];
      {
#pragma unroll
        for (int j = 0; j < kLanePaths; ++j) cont[j] = q[j][0] * c[0];
#pragma unroll
        for (int k = 1; k < KP; ++k) {
#pragma unroll
          for (int j = 0; j < kLanePaths; ++j) cont[j] = __fma_rn(q[j][k], c[k], cont[j]);
        }
      }
#pragma unroll
      for (int j = 0; j < kLanePaths; ++j) {
        const double pay = __fma_rn(price[j], f.price_coef, f.offset);
        const double tt = pay + cont[j];                                       // what the argmax is taken over (LSV.cs:317-327)
        const double w = pay + real[j];
        if (d == 0 || tt > best[j]) {  // first maximum wins (SH.cs:206-221)
          best[j] = tt;
          v[j] = w;
        }
      }
    }
  } else {
    double best[kLanePaths] = {0.0, 0.0, 0.0, 0.0}, adj[kLanePaths] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      const double2 lo01 = *reinterpret_cast<const double2 *>(my_rows + t->lower_row[d]);
      const double2 lo23 = *reinterpret_cast<const double2 *>(my_rows + t->lower_row[d] + kPairGap * 8);
      const double2 up01 = *reinterpret_cast<const double2 *>(my_rows + t->upper_row[d]);
      const double2 up23 = *reinterpret_cast<const double2 *>(my_rows + t->upper_row[d] + kPairGap * 8);
      evaluate_decision<KP>(d == 0, t->a[d], t->upper_w[d], t->c[d], lo01, up01, q[0], q[1], price[0], price[1], df, t->inv_cost,
                            best[0], best[1], adj[0], adj[1]);
      evaluate_decision<KP>(d == 0, t->a[d], t->upper_w[d], t->c[d], lo23, up23, q[2], q[3], price[2], price[3], df, t->inv_cost,
                            best[2], best[3], adj[2], adj[3]);
    }
#pragma unroll
    for (int j = 0; j < kLanePaths; ++j) v[j] = best[j] + adj[j];  // LSV.cs:333
  }
}

template <int KP, bool FUSED>
__global__ void __launch_bounds__(kRingThreads, LSMC_RING_MIN_CTAS) backward_sweep_ring_kernel(BackwardStepArgs a) {
  constexpr int KC = (KP + 1) & ~1;
  constexpr int MB = KP <= 8 ? 1 : 2;
  constexpr int ACC = 8 * MB;
  constexpr int CH = kSweepTilePaths / 4 / kRingWarps;  // k-chunks of 4 paths per warp in the regression GEMM (8)
  using Table = PointTable<KC>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Table *s_tab = reinterpret_cast<Table *>(smem_raw);                        // [kRingTabRing]
  double *s_v = reinterpret_cast<double *>(s_tab + kRingTabRing);            // [kVBufs][kBlockG][kVPitch]
  double *s_acc = s_v + kVBufs * kBlockG * kVPitch;                          // [kAccBufs][kRingWarps][ACC][kBlockG]
  char *s_rows = reinterpret_cast<char *>(s_acc + kAccBufs * kRingWarps * ACC * kBlockG);  // [R][kRowBytes]
  const int R = a.ring_rows;

  const int K = a.K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;  // warp = grid slot
  const int64_t tile = blockIdx.y;
  // the lane's paths: the pair (2 lane, 2 lane + 1) of the tile and the pair 64 columns further, so that every 16-byte
  // access of a warp is one contiguous 512-byte run (no shared-memory bank conflicts, fully coalesced global accesses)
  const int64_t n0 = tile * kSweepTilePaths + 2 * lane;
  const bool validA = n0 < a.N, validB = n0 + kPairGap < a.N;
  // Columns in [N, pitch) are padding of the row buffers: zero in the basis rows and in the terminal values, hence finite
  // in every row.  Lanes that own them compute on them like on any other column; a pair that straddles N is stored (and
  // stays finite), a pair entirely beyond N is not.  Padding never reaches the regression sums: the q_prev fragments are
  // zero beyond N.
  const int64_t c0 = n0;

  const int total_blocks = (a.G + kBlockG - 1) / kBlockG;
  const int bbeg = blockIdx.x * a.subtiles_per_cta;
  const int nblocks = min(bbeg + a.subtiles_per_cta, total_blocks) - bbeg;
  if (nblocks <= 0) return;
  const int gbeg = bbeg * kBlockG;
  const int glimit = min(a.G, (bbeg + nblocks) * kBlockG) - gbeg;

  const Table *tab = reinterpret_cast<const Table *>(a.point_tables) + gbeg;
  const uint32_t pitch_bytes = (uint32_t)a.pitch * 8u;
  const char *tile_rows = reinterpret_cast<const char *>(a.v_next + tile * kSweepTilePaths);
  // this CTA's row windows, staged in shared memory: the thread that issues a generation must not wait on global loads while
  // the other warps wait for it at the block barrier
  SweepBlock *blk = reinterpret_cast<SweepBlock *>(s_rows + (size_t)R * kRowBytes);  // [nblocks]
  for (int i = threadIdx.x; i < nblocks; i += kRingThreads) blk[i] = a.sweep_blocks[bbeg + i];
  __shared__ __align__(8) uint64_t s_mbar[2];    // generation j landed (bulk copies, by byte count)
  __shared__ __align__(8) uint64_t s_eval[2];    // every warp has evaluated block j (values published, ring rows released)
  __shared__ __align__(8) uint64_t s_gemm[2];    // every warp has written its regression-sum slice of block j

  // Generation j = everything block j needs that is not resident yet: the rows its window adds to the ring and, every other
  // block, the next sub-tile of point records.  One elected thread posts the byte count on mbarrier j & 1 and issues the bulk
  // copies; everybody waits on that barrier's phase (j >> 1) & 1 before starting block j.
  auto issue_generation = [&](int j) {  // one whole warp: lane 0 posts the byte count, then every lane issues one row
    uint64_t *bar = &s_mbar[j & 1];
    int first_row, first_slot, count;
    if (j == 0) {
      first_row = blk[0].lo, first_slot = blk[0].lo_slot, count = blk[0].hi - blk[0].lo + 1;
    } else {
      const SweepBlock cur = blk[j - 1], nxt = blk[j];
      first_row = cur.hi + 1, first_slot = cur.hi_slot + 1, count = nxt.hi - cur.hi;
    }
    const int sub = j >> 1;  // records of sub-tile `sub` travel with its first block's generation, one block ahead of its use
    const int rec_first = sub * kRingTabSub;
    const int rec_count = (j & 1) == 0 ? max(0, min(kRingTabSub, glimit - rec_first)) : 0;
    if (lane == 0) {
      mbar_expect_tx(bar, (uint32_t)count * kRowBytes + (uint32_t)rec_count * (uint32_t)sizeof(Table));
      if (rec_count > 0) bulk_load(s_tab + (sub & 1) * kRingTabSub, tab + rec_first, (uint32_t)rec_count * (uint32_t)sizeof(Table), bar);
    }
    __syncwarp();
#pragma unroll 1
    for (int i = lane; i < count; i += 32) {
      int slot_i = first_slot + i;
      if (slot_i >= R) slot_i -= R;
      bulk_load(s_rows + (size_t)slot_i * kRowBytes, row_address(tile_rows, (uint32_t)(first_row + i), pitch_bytes), kRowBytes, bar);
    }
    if (LSMC_RING_PF_AHEAD > 0 && j + LSMC_RING_PF_AHEAD < nblocks) {  // ask L2 for the rows a few generations further on
      const int pf_first = blk[j + LSMC_RING_PF_AHEAD - 1].hi + 1, pf_count = blk[j + LSMC_RING_PF_AHEAD].hi - blk[j + LSMC_RING_PF_AHEAD - 1].hi;
#pragma unroll 1
      for (int i = lane; i < pf_count; i += 32) bulk_prefetch_l2(row_address(tile_rows, (uint32_t)(pf_first + i), pitch_bytes), kRowBytes);
    }
  };
  if (threadIdx.x == 0) {
    mbar_init(&s_mbar[0], 1);
    mbar_init(&s_mbar[1], 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_eval[i], kRingWarps);
      mbar_init(&s_gemm[i], kRingWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 0) {
    issue_generation(0);
    if (nblocks > 1) issue_generation(1);  // its records are the second sub-tile only from block 2 on: (1 & 1) != 0, rows only
  }

  double price[kLanePaths];
#pragma unroll
  for (int j = 0; j < kLanePaths; ++j)
    price[j] = a.spot ? a.spot[min(c0 + (j & 1) + (j >> 1) * kPairGap, a.N - 1)] : a.const_price;
  // orthonormal basis rows of this period for the lane's four paths (continuation value = q . c)
  double q[kLanePaths][KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) {
#pragma unroll
    for (int j = 0; j < kLanePaths; ++j) q[j][k] = 0.0;
    if (k < K && a.q_this) {
      const double2 t01 = *reinterpret_cast<const double2 *>(a.q_this + (size_t)k * a.pitch + c0);
      const double2 t23 = *reinterpret_cast<const double2 *>(a.q_this + (size_t)k * a.pitch + c0 + kPairGap);
      q[0][k] = t01.x;
      q[1][k] = t01.y;
      q[2][k] = t23.x;
      q[3][k] = t23.y;
    }
  }
  if (!a.q_this) {
#pragma unroll
    for (int j = 0; j < kLanePaths; ++j) q[j][0] = 1.0;  // mean mode: continuation value = gamma[g][0]
  }

  // A operand of the regression GEMM for this warp's 32-path slice of the CTA: q_prev[k][slice], fragment layout of
  // mma.m8n8k4 (row = lane / 4 -> k, col = lane % 4 -> path inside a chunk of 4).  Paths beyond N contribute zero.
  const int row = lane >> 2, col = lane & 3;
  double afrag[MB][CH];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      const int k = mb * 8 + row;
      const int64_t p = tile * kSweepTilePaths + 4 * CH * warp + 4 * j + col;
      double x = 0.0;
      if (a.accumulate && k < K && p < a.N) x = a.q_prev ? a.q_prev[(size_t)k * a.pitch + p] : (k == 0 ? 1.0 : 0.0);
      afrag[mb][j] = x;
    }
  }

  const double df = a.settle_df;
  const int last = glimit - 1;
  char *vout = reinterpret_cast<char *>(a.v_this + (size_t)gbeg * a.pitch + n0) + (uint64_t)warp * pitch_bytes;
  const char *my_rows = s_rows + lane * 16;                     // this lane's first pair of columns in every ring row
  double *my_v = s_v + (size_t)warp * kVPitch + 2 * lane;       // + buffer; second point kSlots rows further
  const double *my_b = s_v + (size_t)row * kVPitch + 4 * CH * warp + col;      // B fragments of this warp's slice
  double *my_c = s_acc + (size_t)warp * ACC * kBlockG + row * kBlockG + 2 * col;  // C fragment: [k][g], two adjacent g per lane

  auto combine = [&](int b) {  // partial sums of block b: the warps' slices added in warp order; every warp takes a quarter
    constexpr int kPerWarp = kBlockG * ACC / kRingWarps;  // 16 (or 32) items per warp
    if (lane < kPerWarp) {
      const int item = warp * kPerWarp + lane;
      const int k = item / kBlockG, g8 = item - k * kBlockG;
      const int g = b * kBlockG + g8;
      if (g < glimit) {
        const double *src = s_acc + (size_t)(b % kAccBufs) * kRingWarps * ACC * kBlockG + item;
        double sum = src[0];
#pragma unroll
        for (int w = 1; w < kRingWarps; ++w) sum += src[(size_t)w * ACC * kBlockG];
        a.partials[((size_t)(gbeg + g) * gridDim.y + tile) * ACC + k] = sum;
      }
    }
  };

  // The four warps are NOT kept in lock step: instead of a block barrier each hand-off is an mbarrier that the producer
  // warps arrive on (lane 0, after __syncwarp) and only the consumer of that hand-off waits on, so a warp may run up to a
  // block ahead of the slowest one.  Buffer counts follow from that skew: a warp that has seen s_eval[j] complete knows every
  // warp has finished iteration j-1 entirely.
  //   s_eval[j]: block j evaluated by all  -> the GEMM of block j may read the tile; generation j+2 may overwrite the ring
  //   s_gemm[j]: slices of block j written -> they may be combined
  auto warp_arrive = [&](uint64_t *bar) {
    __syncwarp();
    if (lane == 0) mbar_arrive(bar);
  };
  auto gemm = [&](int bb) {  // C[k][g] = sum over the warp's 32 paths of q_prev[k][n] * V[g][n] for block bb
    const double *vb = my_b + (bb % kVBufs) * (kBlockG * kVPitch);
    double *dst = my_c + (bb % kAccBufs) * (kRingWarps * ACC * kBlockG);
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) {
      double acc0 = 0.0, acc1 = 0.0;
#pragma unroll
      for (int j = 0; j < CH; ++j) dmma_m8n8k4(acc0, acc1, afrag[mb][j], vb[4 * j]);
      *reinterpret_cast<double2 *>(dst + mb * 8 * kBlockG) = make_double2(acc0, acc1);
    }
  };

  for (int b = 0; b < nblocks; ++b) {
    // generation b+1 overwrites ring slots and records that block b-1 was the last to read (generation 1: prologue)
    if (b > 0 && b + 1 < nblocks && warp == (b & 3)) {
      mbar_wait(&s_eval[(b - 1) & 1], ((b - 1) >> 1) & 1);
      issue_generation(b + 1);
    }
    mbar_wait(&s_mbar[b & 1], (b >> 1) & 1);  // this block's rows and records have landed
    const int gA = b * kBlockG + warp, gB = gA + kSlots;
    const Table *tA = s_tab + (min(gA, last) & (kRingTabRing - 1));
    const Table *tB = s_tab + (min(gB, last) & (kRingTabRing - 1));
    double *vt = my_v + (b % kVBufs) * (kBlockG * kVPitch);
    {  // first point: evaluate, store, publish to the regression tile (its registers are free before the second point starts)
      double v[kLanePaths];
      evaluate4_ring<KP, KC, FUSED>(tA, my_rows, q, price, df, v);
      if (gA <= last) {
        if (validA) *reinterpret_cast<double2 *>(vout) = make_double2(v[0], v[1]);
        if (validB) *reinterpret_cast<double2 *>(vout + kPairGap * 8) = make_double2(v[2], v[3]);
      }
      if (a.accumulate) {
        *reinterpret_cast<double2 *>(vt) = make_double2(v[0], v[1]);
        *reinterpret_cast<double2 *>(vt + kPairGap) = make_double2(v[2], v[3]);
      }
    }
    {  // second point
      double v[kLanePaths];
      evaluate4_ring<KP, KC, FUSED>(tB, my_rows, q, price, df, v);
      if (gB <= last) {
        char *vo = vout + (uint64_t)kSlots * pitch_bytes;
        if (validA) *reinterpret_cast<double2 *>(vo) = make_double2(v[0], v[1]);
        if (validB) *reinterpret_cast<double2 *>(vo + kPairGap * 8) = make_double2(v[2], v[3]);
      }
      if (a.accumulate) {
        *reinterpret_cast<double2 *>(vt + kSlots * kVPitch) = make_double2(v[0], v[1]);
        *reinterpret_cast<double2 *>(vt + kSlots * kVPitch + kPairGap) = make_double2(v[2], v[3]);
      }
    }
    vout += (uint64_t)kBlockG * pitch_bytes;
    warp_arrive(&s_eval[b & 1]);
    if (a.accumulate) {
      if (b >= 1) {  // the regression sums run one block behind the evaluation
        mbar_wait(&s_eval[(b - 1) & 1], ((b - 1) >> 1) & 1);
        gemm(b - 1);
        warp_arrive(&s_gemm[(b - 1) & 1]);
      }
      if (b >= 2) {
        mbar_wait(&s_gemm[(b - 2) & 1], ((b - 2) >> 1) & 1);
        combine(b - 2);
      }
    }
  }
  if (a.accumulate) {  // drain
    const int bl = nblocks - 1;
    mbar_wait(&s_eval[bl & 1], (bl >> 1) & 1);
    gemm(bl);
    warp_arrive(&s_gemm[bl & 1]);
    if (bl >= 1) {
      mbar_wait(&s_gemm[(bl - 1) & 1], ((bl - 1) >> 1) & 1);
      combine(bl - 1);
    }
    mbar_wait(&s_gemm[bl & 1], (bl >> 1) & 1);
    combine(bl);
  }
}

template <int KP>
void launch_sweep_ring_for_k(const BackwardStepArgs &a, dim3 grid, size_t smem, cudaStream_t stream) {
  if (a.fused_payoff) {
    ensure_dynamic_smem<backward_sweep_ring_kernel<KP, true>>(smem);
    backward_sweep_ring_kernel<KP, true><<<grid, kRingThreads, smem, stream>>>(a);
  } else {
    ensure_dynamic_smem<backward_sweep_ring_kernel<KP, false>>(smem);
    backward_sweep_ring_kernel<KP, false><<<grid, kRingThreads, smem, stream>>>(a);
  }
}

template <int KP>
void launch_sweep_for_k(const BackwardStepArgs &a, dim3 grid, size_t smem, cudaStream_t stream) {
  if (a.fused_payoff) {
    ensure_dynamic_smem<backward_sweep_kernel<KP, true>>(smem);
    backward_sweep_kernel<KP, true><<<grid, kSweepThreads, smem, stream>>>(a);
  } else {
    ensure_dynamic_smem<backward_sweep_kernel<KP, false>>(smem);
    backward_sweep_kernel<KP, false><<<grid, kSweepThreads, smem, stream>>>(a);
  }
}

template <int KC>
void launch_fold_for_kc(const FoldArgs &f, cudaStream_t stream) {
  fold_tables_kernel<KC><<<(fold_item_count(f) + 127) / 128, 128, 0, stream>>>(f);
}

int coef_width(int KP) { return (KP + 1) & ~1; }

}  // namespace

size_t point_table_bytes(int K) {
  switch (coef_width(padded_basis(K))) {
    case 2: return sizeof(PointTable<2>);
    case 4: return sizeof(PointTable<4>);
    case 6: return sizeof(PointTable<6>);
    case 8: return sizeof(PointTable<8>);
    case 12: return sizeof(PointTable<12>);
    default: return sizeof(PointTable<16>);
  }
}

size_t backward_sweep_smem_bytes(int K) {
  const int KP = padded_basis(K);
  const int ACC = KP <= 8 ? 8 : 16;
  return point_table_bytes(K) * kTabRing + sizeof(double) * (2 * kBlockG * kVPitch + 2 * kSweepWarps * kBlockG * ACC);
}

size_t backward_sweep_ring_smem_bytes(int K, int ring_rows, int blocks_per_cta) {
  const int KP = padded_basis(K);
  const int ACC = KP <= 8 ? 8 : 16;
  return point_table_bytes(K) * kRingTabRing + sizeof(double) * (kVBufs * kBlockG * kVPitch + kAccBufs * kRingWarps * kBlockG * ACC) +
         (size_t)ring_rows * kRowBytes + (size_t)blocks_per_cta * sizeof(SweepBlock);
}

template <int KC>
void launch_reduce_fold_for_kc(const ReduceArgs &r, const FoldArgs &f, unsigned int *counter, cudaStream_t stream) {
  reduce_fold_kernel<KC><<<r.G, kReduceThreads, 0, stream>>>(r, f, counter);
}

void launch_reduce_fold(const ReduceArgs &r, const FoldArgs &f, unsigned int *counter, cudaStream_t stream) {
  switch (coef_width(padded_basis(f.K))) {
    case 2: launch_reduce_fold_for_kc<2>(r, f, counter, stream); break;
    case 4: launch_reduce_fold_for_kc<4>(r, f, counter, stream); break;
    case 6: launch_reduce_fold_for_kc<6>(r, f, counter, stream); break;
    case 8: launch_reduce_fold_for_kc<8>(r, f, counter, stream); break;
    case 12: launch_reduce_fold_for_kc<12>(r, f, counter, stream); break;
    default: launch_reduce_fold_for_kc<16>(r, f, counter, stream); break;
  }
}

template <int KC>
void launch_exchange_fold_for_kc(const ReduceArgs &r, const FoldArgs &f, const XchgArgs &x, unsigned int *counter, cudaStream_t stream) {
  exchange_fold_kernel<KC><<<r.G, kReduceThreads, 0, stream>>>(r, f, x, counter);
}

void launch_exchange_fold(const ReduceArgs &r, const FoldArgs &f, const XchgArgs &x, unsigned int *counter, cudaStream_t stream) {
  switch (coef_width(padded_basis(r.K))) {
    case 2: launch_exchange_fold_for_kc<2>(r, f, x, counter, stream); break;
    case 4: launch_exchange_fold_for_kc<4>(r, f, x, counter, stream); break;
    case 6: launch_exchange_fold_for_kc<6>(r, f, x, counter, stream); break;
    case 8: launch_exchange_fold_for_kc<8>(r, f, x, counter, stream); break;
    case 12: launch_exchange_fold_for_kc<12>(r, f, x, counter, stream); break;
    default: launch_exchange_fold_for_kc<16>(r, f, x, counter, stream); break;
  }
}

void launch_fold_tables(const FoldArgs &f, cudaStream_t stream) {
  switch (coef_width(padded_basis(f.K))) {
    case 2: launch_fold_for_kc<2>(f, stream); break;
    case 4: launch_fold_for_kc<4>(f, stream); break;
    case 6: launch_fold_for_kc<6>(f, stream); break;
    case 8: launch_fold_for_kc<8>(f, stream); break;
    case 12: launch_fold_for_kc<12>(f, stream); break;
    default: launch_fold_for_kc<16>(f, stream); break;
  }
}

void launch_backward_sweep(const BackwardStepArgs &a, int gsplit, int64_t num_tiles, cudaStream_t stream) {
  const dim3 grid(gsplit, (unsigned)num_tiles);
  if (a.ring_rows > 0) {
    const size_t smem = backward_sweep_ring_smem_bytes(a.K, a.ring_rows, a.subtiles_per_cta);
    switch (padded_basis(a.K)) {
      case 1: launch_sweep_ring_for_k<1>(a, grid, smem, stream); break;
      case 2: launch_sweep_ring_for_k<2>(a, grid, smem, stream); break;
      case 3: launch_sweep_ring_for_k<3>(a, grid, smem, stream); break;
      case 4: launch_sweep_ring_for_k<4>(a, grid, smem, stream); break;
      case 5: launch_sweep_ring_for_k<5>(a, grid, smem, stream); break;
      case 6: launch_sweep_ring_for_k<6>(a, grid, smem, stream); break;
      case 7: launch_sweep_ring_for_k<7>(a, grid, smem, stream); break;
      case 8: launch_sweep_ring_for_k<8>(a, grid, smem, stream); break;
      case 12: launch_sweep_ring_for_k<12>(a, grid, smem, stream); break;
      default: launch_sweep_ring_for_k<16>(a, grid, smem, stream); break;
    }
    return;
  }
  const size_t smem = backward_sweep_smem_bytes(a.K);
  switch (padded_basis(a.K)) {
    case 1: launch_sweep_for_k<1>(a, grid, smem, stream); break;
    case 2: launch_sweep_for_k<2>(a, grid, smem, stream); break;
    case 3: launch_sweep_for_k<3>(a, grid, smem, stream); break;
    case 4: launch_sweep_for_k<4>(a, grid, smem, stream); break;
    case 5: launch_sweep_for_k<5>(a, grid, smem, stream); break;
    case 6: launch_sweep_for_k<6>(a, grid, smem, stream); break;
    case 7: launch_sweep_for_k<7>(a, grid, smem, stream); break;
    case 8: launch_sweep_for_k<8>(a, grid, smem, stream); break;
    case 12: launch_sweep_for_k<12>(a, grid, smem, stream); break;
    default: launch_sweep_for_k<16>(a, grid, smem, stream); break;
  }
}

}  // namespace lsmc
