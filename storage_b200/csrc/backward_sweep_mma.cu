// backward_sweep_mma.cu — the backward-induction step (LSV.cs:240-338) for three-decision tables in the fused-payoff form,
// with the payoff + regressed continuation value of every decision formed on the FP64 tensor path (mma.sync.m8n8k4.f64).
//
// Why.  The DFMA sweep kernel (backward_sweep.cu) is bound by instruction issue and by the FP64 pipe together: per 128
// path x grid-point evaluations a warp issues ~340 instructions of which ~170 are FP64, and the FP64 pipe of B200 accepts one
// warp-wide DFMA every ~2.2 cycles per scheduler (profiles/micro/fp64_pipes.cu).  DMMA runs on the same pipe at the same
// FLOP rate (one m8n8k4 every 16 cycles per scheduler) but moves 256 FMAs per issued instruction instead of 32 and takes
// its coefficient operand as a per-lane fragment instead of broadcast shared-memory loads.  The evaluation is therefore
// re-cast as small GEMMs:
//
//   T_d[g][n] = offset_d(g) + sum_k c_d(g)[k] * qx[n][k]      d = 0,1,2     (8 grid points x 8 paths x K+1)
//
// where qx = (orthonormal basis row of the path, spot price of the path) and c_d(g) = (regression coefficients with the
// next-grid interpolation folded in, (-v_d - consumed_d) * DF): T_d is the total the argmax is taken over (immediate cash
// flow + regressed continuation value, LSV.cs:317-327).  The fragment layout puts grid point r = lane / 4 and the adjacent
// paths 2c, 2c+1 (c = lane % 4) of an 8-path group in one lane, which is also what the rest of the step wants: the rows of
// V_{p+1} are read as 16-byte pairs, V_p is stored as 16-byte pairs, and the values feed the regression GEMM of the next
// backward step (Q_{p-1}^T V_p, also DMMA) straight from registers as its B fragments — no shared-memory exchange.
//
// Work layout.  A CTA owns 128 paths (4 warps x 32 paths = 4 groups of 8) and sweeps its range of the inventory grid upwards in
// blocks of 8 grid points; every warp evaluates all 8 grid points of a block for its own paths.  The rows of V_{p+1} a block
// reads live in a shared-memory ring filled by bulk asynchronous copies (TMA 1-D) kMmaLookAhead blocks ahead, together with
// the block's record (mma_records.cuh); hand-offs are mbarriers, there is no CTA barrier in the loop.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include "backward_common.cuh"
#include "device_types.hpp"
#include "kernels.hpp"
#include "mma_records.cuh"
#include "sweep_common.cuh"

#ifndef LSMC_MMA_MIN_CTAS
#define LSMC_MMA_MIN_CTAS 3
#endif
#ifndef LSMC_MMA_WINNER_ONLY
#define LSMC_MMA_WINNER_ONLY 1  // 0: form the carried value for all three decisions and select the value (round 2's first form)
#endif
#ifndef LSMC_MMA_MEMONLY
#define LSMC_MMA_MEMONLY 0  // diagnostic build only (profiles/ab_library.sh): the kernel without its mma work; results are wrong
#endif

namespace lsmc {

namespace {

#ifndef LSMC_MMA_GROUPS
#define LSMC_MMA_GROUPS 4
#endif
constexpr int kMmaGroups = LSMC_MMA_GROUPS;       // 8-path groups per warp
constexpr int kMmaWarps = kSweepTilePaths / (8 * kMmaGroups);
constexpr int kMmaThreads = 32 * kMmaWarps;
constexpr int kMmaBlockG = 8;                     // grid points per block = m of the m8n8k4 tile
constexpr int kMmaGens = kMmaLookAhead + 1 + kMmaSlack;  // generations (rows + record of one block) in flight, in use or not yet released
constexpr int kMmaRowBytes = kSweepTilePaths * 8;
#ifndef LSMC_MMA_ACC_LAG
#define LSMC_MMA_ACC_LAG 2
#endif
constexpr int kAccLag = LSMC_MMA_ACC_LAG;         // blocks between a warp writing its regression-sum slice and the slices being combined
constexpr int kAccBufs = 2 * kAccLag;             // slice buffers: a warp may be kAccLag blocks ahead of the slowest one
constexpr int kAccBars = kAccLag < 2 ? 2 : kAccLag;  // hand-off barriers: block b + kAccLag must not share one with a block still open
static_assert(kMmaWarps * kMmaGroups * 8 == kSweepTilePaths && (kMmaWarps & (kMmaWarps - 1)) == 0, "warps x groups x 8 paths = one 128-path tile");

template <int KSTEPS, int MB>
__global__ void __launch_bounds__(kMmaThreads, LSMC_MMA_MIN_CTAS) backward_sweep_mma_kernel(BackwardStepArgs a) {
  using Rec = MmaBlockRecord<KSTEPS>;
  constexpr int ACC = 8 * MB;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int R = a.ring_rows;
  char *s_rows = reinterpret_cast<char *>(smem_raw);                                      // [R][kMmaRowPitchBytes]
  Rec *s_rec = reinterpret_cast<Rec *>(s_rows + (size_t)R * kMmaRowPitchBytes);            // [kMmaGens]
  double *s_acc = reinterpret_cast<double *>(s_rec + kMmaGens);                            // [kAccBufs][kMmaWarps][ACC][8]
  SweepBlock *blk = reinterpret_cast<SweepBlock *>(s_acc + kAccBufs * kMmaWarps * ACC * kMmaBlockG);  // [nblocks]
  __shared__ __align__(8) uint64_t s_mbar[kMmaGens];  // generation j landed (bulk copies, by byte count)
  __shared__ __align__(8) uint64_t s_eval[kMmaGens];  // every warp has evaluated block j: its ring rows and record may be overwritten
  __shared__ __align__(8) uint64_t s_gemm[kAccBars];  // every warp has written its regression-sum slice of block j

  const int K = a.K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r = lane >> 2, c = lane & 3;

  // Work split.  The first a.whole_tiles CTAs take one whole 128-path tile each (all blocks of 8 grid points); the hardware
  // hands them to SM slots as slots free up, which balances the SMs dynamically.  The remaining tiles — the last, partial
  // wave — are cut into a.tail_split pieces each, so that the grid ends on many short CTAs instead of a few long ones (977
  // tiles on 444 slots ran as 2.2 waves and paid for 3).
  const int total_blocks = (a.G + kMmaBlockG - 1) / kMmaBlockG;
  int64_t tile = blockIdx.x;
  int bbeg = 0, nblocks = total_blocks;
  if (tile >= a.whole_tiles) {
    const int64_t i = tile - a.whole_tiles;
    const int piece = (int)(i % a.tail_split);
    tile = a.whole_tiles + i / a.tail_split;
    bbeg = total_blocks * piece / a.tail_split;
    nblocks = total_blocks * (piece + 1) / a.tail_split - bbeg;
  }
  if (nblocks <= 0) return;
  const int gbeg = bbeg * kMmaBlockG;
  const int glimit = min(a.G, (bbeg + nblocks) * kMmaBlockG) - gbeg;  // grid points of this CTA, addressed 0..glimit-1 below

  const Rec *recs = reinterpret_cast<const Rec *>(a.point_tables) + bbeg;
  const uint32_t pitch_bytes = (uint32_t)a.pitch * 8u;
  const char *tile_rows = reinterpret_cast<const char *>(a.v_next + tile * kSweepTilePaths);
  for (int i = threadIdx.x; i < nblocks; i += kMmaThreads) blk[i] = a.sweep_blocks[bbeg + i];

  // Generation j = what block j needs that is not resident yet: the rows its window adds to the ring and its record.
  auto issue_record = [&](int j) {  // the block's record (written by the kernel that precedes this one in the stream)
    if (lane == 0) bulk_load(s_rec + j % kMmaGens, recs + j, (uint32_t)sizeof(Rec), &s_mbar[j % kMmaGens]);
  };
  auto issue_rows = [&](int j) {  // one whole warp: lane 0 posts the generation's byte count, then every lane issues one row
    uint64_t *bar = &s_mbar[j % kMmaGens];
    int first_row, first_slot, count;
    if (j == 0) {
      first_row = blk[0].lo, first_slot = blk[0].lo_slot, count = blk[0].hi - blk[0].lo + 1;
    } else {
      const SweepBlock cur = blk[j - 1], nxt = blk[j];
      first_row = cur.hi + 1, first_slot = cur.hi_slot + 1, count = nxt.hi - cur.hi;
    }
    if (lane == 0) mbar_expect_tx(bar, (uint32_t)count * kMmaRowBytes + (uint32_t)sizeof(Rec));
    __syncwarp();
#pragma unroll 1
    for (int i = lane; i < count; i += 32) {
      int slot_i = first_slot + i;
      if (slot_i >= R) slot_i -= R;
      bulk_load(s_rows + (size_t)slot_i * kMmaRowPitchBytes, row_address(tile_rows, (uint32_t)(first_row + i), pitch_bytes), kMmaRowBytes, bar);
    }
  };
  if (threadIdx.x == 0) {
    for (int i = 0; i < kMmaGens; ++i) {
      mbar_init(&s_mbar[i], 1);
      mbar_init(&s_eval[i], kMmaWarps);
    }
    for (int i = 0; i < kAccBars; ++i) mbar_init(&s_gemm[i], kMmaWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // Programmatic dependent launch: this grid may start while the per-step reduction / exchange + fold kernel before it is
  // still running (that kernel lets its dependents go at once).  Everything up to here, the first generations' rows (V_{p+1}
  // is older than that kernel) and the per-path operand loads below overlap its tail — its last CTA folds the records alone
  // while the rest of the GPU would idle; only the records must wait for it.  Launched without the attribute the wait is a no-op.
  if (warp == 0) {
    for (int j = 0; j < min(kMmaLookAhead + 1, nblocks); ++j) issue_rows(j);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for (int j = 0; j < min(kMmaLookAhead + 1, nblocks); ++j) issue_record(j);
  }

  // ---- per-path operands, loaded once
  // Columns in [N, pitch) are padding of the row buffers: zero in the basis rows and in the terminal values, hence finite in
  // every row.  Lanes that own them compute on them like on any other column; a pair that straddles N is stored (and stays
  // finite), a pair entirely beyond N is not.  Padding never reaches the regression sums: the q_prev fragments are zero there.
  const int64_t warp_path0 = tile * kSweepTilePaths + 8 * kMmaGroups * warp;
  // B fragments of qx = (basis row, spot price): lane (r, c) supplies coefficient 4 s + c of path 8 j + r of the warp
  double qb[kMmaGroups][KSTEPS];
#pragma unroll
  for (int j = 0; j < kMmaGroups; ++j) {
    const int64_t p = warp_path0 + 8 * j + r;
#pragma unroll
    for (int s = 0; s < KSTEPS; ++s) {
      const int k = 4 * s + c;
      double x = 0.0;
      if (k < K)
        x = a.q_this ? a.q_this[(size_t)k * a.pitch + p] : (k == 0 ? 1.0 : 0.0);  // mean mode: continuation value = gamma[g][0]
      else if (k == K)
        x = a.spot ? a.spot[min(p, a.N - 1)] : a.const_price;
      qb[j][s] = x;
    }
  }
  // the lane's own paths: 8 j + 2 c and + 1 of the warp
  const int64_t own0 = warp_path0 + 2 * c;
  double2 price[kMmaGroups];
#pragma unroll
  for (int j = 0; j < kMmaGroups; ++j) {
    const int64_t p = own0 + 8 * j;
    price[j] = a.spot ? make_double2(a.spot[min(p, a.N - 1)], a.spot[min(p + 1, a.N - 1)]) : make_double2(a.const_price, a.const_price);
  }
  // A fragments of the regression GEMM C[k][g] = sum_n q_prev[k][n] V[g][n]: lane (r, c) supplies basis function r (+ 8 mb) of
  // path 2 c (.x) and 2 c + 1 (.y) of every group; the two halves are the two k4 chunks (even / odd paths) of the group
  double2 qp[kMmaGroups][MB];
#pragma unroll
  for (int j = 0; j < kMmaGroups; ++j) {
    const int64_t p = own0 + 8 * j;
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) {
      const int k = 8 * mb + r;
      double2 x = make_double2(0.0, 0.0);
      if (a.accumulate && k < K) {
        if (a.q_prev)
          x = *reinterpret_cast<const double2 *>(a.q_prev + (size_t)k * a.pitch + p);
        else if (k == 0)
          x = make_double2(1.0, 1.0);  // next regression is the sample mean: plain sum of V
        if (p >= a.N) x.x = 0.0;
        if (p + 1 >= a.N) x.y = 0.0;
      }
      qp[j][mb] = x;
    }
  }

  const int last = glimit - 1;
  const uint32_t own_col_bytes = (uint32_t)(8 * kMmaGroups * warp + 2 * c) * 8u;   // the lane's first pair inside a ring row
  char *vout = reinterpret_cast<char *>(a.v_this + (size_t)(gbeg + r) * a.pitch + own0);
  double *my_acc = s_acc + (size_t)warp * ACC * kMmaBlockG + r * kMmaBlockG + 2 * c;  // C fragment: [k][g], two adjacent g per lane

  auto warp_arrive = [&](uint64_t *bar) {
    __syncwarp();
    if (lane == 0) mbar_arrive(bar);
  };
  auto combine = [&](int b) {  // partial sums of block b: the warps' slices added in warp order; every warp takes a quarter
    constexpr int kPerWarp = kMmaBlockG * ACC / kMmaWarps;  // 16 (or 32) items per warp
    if (lane < kPerWarp) {
      const int item = warp * kPerWarp + lane;
      const int k = item / kMmaBlockG, g8 = item - k * kMmaBlockG;
      const int g = b * kMmaBlockG + g8;
      if (g < glimit) {
        const double *src = s_acc + (size_t)(b % kAccBufs) * kMmaWarps * ACC * kMmaBlockG + item;
        double sum = src[0];
#pragma unroll
        for (int w = 1; w < kMmaWarps; ++w) sum += src[(size_t)w * ACC * kMmaBlockG];
        a.partials[((size_t)(gbeg + g) * a.num_tiles + tile) * ACC + k] = sum;
      }
    }
  };

  for (int b = 0; b < nblocks; ++b) {
    // generation b + look-ahead overwrites ring slots and the record slot that block b - 1 was the last to read.  The issuing
    // warp may spin here (6.6 % of the kernel's stall samples), but it is the warp that is ahead: issuing after its own block
    // instead, with one more block of look-ahead, was measured 5-7 % slower.
    if (b > 0 && b + kMmaLookAhead < nblocks && warp == (b & (kMmaWarps - 1))) {
      if (b - 1 - kMmaSlack >= 0) mbar_wait(&s_eval[(b - 1 - kMmaSlack) % kMmaGens], ((b - 1 - kMmaSlack) / kMmaGens) & 1);
      issue_rows(b + kMmaLookAhead);
      issue_record(b + kMmaLookAhead);
    }
    mbar_wait(&s_mbar[b % kMmaGens], (b / kMmaGens) & 1);  // this block's rows and record have landed
    const Rec *rec = s_rec + b % kMmaGens;
    MmaScalars sc[3];
    uint2 rw[3];
    double af[3][KSTEPS];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      sc[d] = rec->s[d][r];
      rw[d] = rec->rows[d][r];
#pragma unroll
      for (int s = 0; s < KSTEPS; ++s) af[d][s] = rec->a[d][s][lane];
    }
    const bool store_g = b * kMmaBlockG + r <= last;
    double2 racc[MB];
#pragma unroll
    for (int mb = 0; mb < MB; ++mb) racc[mb] = make_double2(0.0, 0.0);
    const char *my_rows = s_rows + own_col_bytes;
#pragma unroll
    for (int j = 0; j < kMmaGroups; ++j) {
      // totals the argmax is taken over, for grid point r and the paths 8 j + 2 c, + 1 (LSV.cs:317-327)
      double t[3][2];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
#if LSMC_MMA_MEMONLY  // diagnostic build: the kernel's data movement without its tensor work (results are wrong)
        t[d][0] = sc[d].offset + af[d][0] * qb[j][0], t[d][1] = sc[d].offset_again + af[d][KSTEPS - 1] * qb[j][KSTEPS - 1];
#else
        dmma_m8n8k4_to(t[d][0], t[d][1], af[d][0], qb[j][0], sc[d].offset, sc[d].offset_again);
#pragma unroll
        for (int s = 1; s < KSTEPS; ++s) dmma_m8n8k4(t[d][0], t[d][1], af[d][s], qb[j][s]);
#endif
      }
#if LSMC_MMA_WINNER_ONLY
      // first maximum wins (SH.cs:206-221); the value carried is formed for the winning decision only: immediate cash flow +
      // realised continuation value (LSV.cs:296-302, 330-335), the weighted average of the two rows of V_{p+1} the decision
      // lands between (an exact grid hit has lower == upper and weights (1, 0): the row itself).  Selecting the winner's
      // scalars and row offsets costs ALU selects but a third of the FP64 operations and of the shared-memory wavefronts of
      // forming the value for all three decisions.
      double2 v;
      {
        const bool a1 = t[1][0] > t[0][0];
        const double b1 = a1 ? t[1][0] : t[0][0];
        const bool a2 = t[2][0] > b1;
        auto pick = [&](auto x0, auto x1, auto x2) { return a2 ? x2 : (a1 ? x1 : x0); };
        const double lo = *reinterpret_cast<const double *>(my_rows + pick(rw[0].x, rw[1].x, rw[2].x) + j * 64);
        const double up = *reinterpret_cast<const double *>(my_rows + pick(rw[0].y, rw[1].y, rw[2].y) + j * 64);
        v.x = __fma_rn(price[j].x, pick(sc[0].price_coef, sc[1].price_coef, sc[2].price_coef),
                       __fma_rn(up, pick(sc[0].upper_w, sc[1].upper_w, sc[2].upper_w),
                                __fma_rn(lo, pick(sc[0].lower_w, sc[1].lower_w, sc[2].lower_w), pick(sc[0].offset, sc[1].offset, sc[2].offset))));
      }
      {
        const bool a1 = t[1][1] > t[0][1];
        const double b1 = a1 ? t[1][1] : t[0][1];
        const bool a2 = t[2][1] > b1;
        auto pick = [&](auto x0, auto x1, auto x2) { return a2 ? x2 : (a1 ? x1 : x0); };
        const double lo = *reinterpret_cast<const double *>(my_rows + pick(rw[0].x, rw[1].x, rw[2].x) + j * 64 + 8);
        const double up = *reinterpret_cast<const double *>(my_rows + pick(rw[0].y, rw[1].y, rw[2].y) + j * 64 + 8);
        v.y = __fma_rn(price[j].y, pick(sc[0].price_coef, sc[1].price_coef, sc[2].price_coef),
                       __fma_rn(up, pick(sc[0].upper_w, sc[1].upper_w, sc[2].upper_w),
                                __fma_rn(lo, pick(sc[0].lower_w, sc[1].lower_w, sc[2].lower_w), pick(sc[0].offset, sc[1].offset, sc[2].offset))));
      }
#else
      // value carried for a decision: immediate cash flow + realised continuation value (LSV.cs:296-302, 330-335), the weighted
      // average of the two rows of V_{p+1}; an exact grid hit has lower == upper and weights (1, 0): the row itself
      double2 w[3];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const double2 lo = *reinterpret_cast<const double2 *>(my_rows + rw[d].x + j * 64);
        const double2 up = *reinterpret_cast<const double2 *>(my_rows + rw[d].y + j * 64);
        w[d].x = __fma_rn(price[j].x, sc[d].price_coef, __fma_rn(up.x, sc[d].upper_w, __fma_rn(lo.x, sc[d].lower_w, sc[d].offset)));
        w[d].y = __fma_rn(price[j].y, sc[d].price_coef, __fma_rn(up.y, sc[d].upper_w, __fma_rn(lo.y, sc[d].lower_w, sc[d].offset)));
      }
      double2 v = w[0];
      double best0 = t[0][0], best1 = t[0][1];
#pragma unroll
      for (int d = 1; d < 3; ++d) {  // first maximum wins (SH.cs:206-221)
        if (t[d][0] > best0) {
          best0 = t[d][0];
          v.x = w[d].x;
        }
        if (t[d][1] > best1) {
          best1 = t[d][1];
          v.y = w[d].y;
        }
      }
#endif
      if (store_g && own0 + 8 * j < a.N) *reinterpret_cast<double2 *>(vout + j * 64) = v;
      if (a.accumulate) {
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) {
#if LSMC_MMA_MEMONLY
          racc[mb].x += qp[j][mb].x * v.x, racc[mb].y += qp[j][mb].y * v.y;
#else
          dmma_m8n8k4(racc[mb].x, racc[mb].y, qp[j][mb].x, v.x);
          dmma_m8n8k4(racc[mb].x, racc[mb].y, qp[j][mb].y, v.y);
#endif
        }
      }
    }
    vout += (uint64_t)kMmaBlockG * pitch_bytes;
    warp_arrive(&s_eval[b % kMmaGens]);
    if (a.accumulate) {
      // The slices of a block are combined kAccLag blocks later: by then the other warps have arrived and the wait rarely stalls
      // (measured: lag 2 is 1-2 % faster than lag 1 and than lag 3).  A warp that has seen s_gemm of block b - kAccLag complete
      // knows every warp has combined block b - 2 kAccLag: that buffer is free.
      if (b >= kAccLag) {
        mbar_wait(&s_gemm[(b - kAccLag) % kAccBars], ((b - kAccLag) / kAccBars) & 1);
        combine(b - kAccLag);
      }
      double *dst = my_acc + (size_t)(b % kAccBufs) * kMmaWarps * ACC * kMmaBlockG;
#pragma unroll
      for (int mb = 0; mb < MB; ++mb) *reinterpret_cast<double2 *>(dst + mb * 8 * kMmaBlockG) = racc[mb];
      warp_arrive(&s_gemm[b % kAccBars]);
    }
  }
  if (a.accumulate) {
    for (int bb = max(0, nblocks - kAccLag); bb < nblocks; ++bb) {
      mbar_wait(&s_gemm[bb % kAccBars], (bb / kAccBars) & 1);
      combine(bb);
    }
  }
}

template <int KSTEPS, int MB>
void launch_mma(const BackwardStepArgs &a, dim3 grid, size_t smem, cudaStream_t stream) {
  ensure_dynamic_smem<backward_sweep_mma_kernel<KSTEPS, MB>>(smem);
  static const bool pdl = [] { const char *e = std::getenv("LSMC_PDL"); return !(e && e[0] == '0'); }();  // test hook: 0 = plain launches
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kMmaThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, backward_sweep_mma_kernel<KSTEPS, MB>, a);
  const cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess)
    throw std::runtime_error("backward_sweep_mma_kernel<" + std::to_string(KSTEPS) + "," + std::to_string(MB) + "> launch failed (" + cudaGetErrorString(e) +
                             "): grid " + std::to_string(grid.x) + ", dynamic shared memory " + std::to_string(smem) + " B, ring rows " +
                             std::to_string(a.ring_rows) + ", G " + std::to_string(a.G) + ", tiles " + std::to_string((long long)a.num_tiles));
}

}  // namespace

size_t mma_block_record_bytes(int K) { return mma_record_bytes(mma_ksteps(K)); }

int backward_sweep_mma_ctas_per_sm() { return LSMC_MMA_MIN_CTAS; }

size_t backward_sweep_mma_smem_bytes(int K, int ring_rows, int blocks_per_cta) {
  const int acc = K <= 8 ? 8 : 16;
  return (size_t)ring_rows * kMmaRowPitchBytes + kMmaGens * mma_block_record_bytes(K) + sizeof(double) * kAccBufs * kMmaWarps * acc * kMmaBlockG +
         (size_t)blocks_per_cta * sizeof(SweepBlock);
}

void launch_backward_sweep_mma(const BackwardStepArgs &a, cudaStream_t stream) {
  const dim3 grid((unsigned)(a.whole_tiles + (a.num_tiles - a.whole_tiles) * a.tail_split));
  const size_t smem = backward_sweep_mma_smem_bytes(a.K, a.ring_rows, (a.G + kMmaBlockG - 1) / kMmaBlockG);
  const int ks = mma_ksteps(a.K);
  const bool wide = a.K > 8;
  static_assert(sizeof(MmaBlockRecord<2>) == 3 * 2 * 32 * 8 + 3 * 8 * sizeof(MmaScalars) + 3 * 8 * 8, "record layout is what mma_record_bytes() says");
  switch (ks) {
    case 1: launch_mma<1, 1>(a, grid, smem, stream); break;
    case 2: launch_mma<2, 1>(a, grid, smem, stream); break;
    case 3:
      if (wide)
        launch_mma<3, 2>(a, grid, smem, stream);
      else
        launch_mma<3, 1>(a, grid, smem, stream);
      break;
    case 4: launch_mma<4, 2>(a, grid, smem, stream); break;
    default: launch_mma<5, 2>(a, grid, smem, stream); break;
  }
}

}  // namespace lsmc
