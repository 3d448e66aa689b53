// backward_ring.cu — backward-induction step with the rows of v_next staged in shared memory by bulk TMA copies.
//
// One CTA = one tile of 256 paths (128 threads x double2) sweeping a contiguous range of inventory-grid points in
// groups of 8.  The rows of v_next that a group's decisions interpolate between form a window [win_lo, win_hi] that
// slides upwards with the grid (the host checks this; otherwise it launches the direct-load kernel of kernels.cu).  The
// window lives in a ring of R rows x 2 KB in shared memory: while a group is being evaluated, one thread has already
// issued cp.async.bulk copies (completing on an mbarrier) for the rows the next group adds, so every row of v_next is
// fetched from HBM/L2 exactly once per CTA, asynchronously, and the decision loop only issues conflict-free LDS.128.
//
// The decision arithmetic follows LSV.cs:310-336 operation for operation (see kernels.cu for the line-by-line notes);
// the regression sums of the next backward step run on the FP64 tensor path (mma.sync m8n8k4) per warp.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#include "backward_common.cuh"
#include "device_types.hpp"
#include "kernels.hpp"

namespace lsmc {

namespace {

constexpr int kGroup = 8;                         // grid points per ring / mma group
constexpr int kRowBytes = kTilePaths * 8;         // one row of the tile: 2 KB

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_row(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

#ifdef LSMC_RING_CHECK
#define CHECK16(ptr, id)                                                                                              \
  do {                                                                                                                \
    if (((unsigned long long)(ptr)) & 15ull)                                                                          \
      printf("misaligned access id=%d ptr=%p tid=%d blk=(%d,%d)\n", id, (const void *)(ptr), (int)threadIdx.x, (int)blockIdx.x, (int)blockIdx.y); \
  } while (0)
#else
#define CHECK16(ptr, id)
#endif

template <int KP>
__global__ void __launch_bounds__(kThreads, 2) backward_ring_kernel(BackwardStepArgs a) {
  constexpr int ACC = KP <= 8 ? 8 : 16;
  constexpr int KC = (KP + 1) & ~1;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int K = a.K, D = a.D, R = a.ring_rows;
  // dynamic smem: ring [R][256] | decA [8][D] | decB [8][D] | coef [8][D][KC] | vtile [kWarps][8][pitch]
  //               | qtile [kWarps][ACC][pitch] | acc [kWarps][ACC][8] | inv_cost [8] | nd [8] | mbarrier
  double *s_ring = reinterpret_cast<double *>(smem_raw);
  DecisionA *s_a = reinterpret_cast<DecisionA *>(s_ring + (size_t)R * kTilePaths);
  DecisionR *s_b = reinterpret_cast<DecisionR *>(s_a + (size_t)kGroup * D);
  double *s_coef = reinterpret_cast<double *>(s_b + (size_t)kGroup * D);
  double *s_vtile = s_coef + (size_t)kGroup * D * KC;
  double *s_qtile = s_vtile + (size_t)kWarps * 8 * kTilePitch;
  double *s_acc = s_qtile + (size_t)kWarps * ACC * kTilePitch;
  double *s_inv_cost = s_acc + (size_t)kWarps * ACC * kGroup;
  int32_t *s_nd = reinterpret_cast<int32_t *>(s_inv_cost + kGroup);
  uint64_t *s_bar = reinterpret_cast<uint64_t *>(s_nd + kGroup);

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile = blockIdx.y;
  const int64_t n0 = tile * kTilePaths + (int64_t)threadIdx.x * kPathsPerThread;
  const bool valid0 = n0 < a.N, valid1 = n0 + 1 < a.N;
  const int64_t c0 = valid0 ? n0 : ((a.N - 1) & ~int64_t(1));
  const int64_t c1 = valid1 ? n0 + 1 : a.N - 1;

  double price0, price1;
  if (a.spot) {
    price0 = a.spot[valid0 ? n0 : a.N - 1];
    price1 = a.spot[c1];
  } else {
    price0 = price1 = a.const_price;
  }
  double q0[KP], q1[KP];
#pragma unroll
  for (int k = 0; k < KP; ++k) {
    q0[k] = q1[k] = 0.0;
    if (k < K && a.q_this) {
      CHECK16(a.q_this + (size_t)k * a.pitch + c0, 1);
      const double2 t = *reinterpret_cast<const double2 *>(a.q_this + (size_t)k * a.pitch + c0);
      q0[k] = t.x;
      q1[k] = t.y;
    }
  }
  if (!a.q_this) q0[0] = q1[0] = 1.0;  // mean mode: continuation value = gamma[g][0]

  double *my_q = s_qtile + (size_t)warp * ACC * kTilePitch;
  double *my_v = s_vtile + (size_t)warp * 8 * kTilePitch;
  if (a.accumulate) {
#pragma unroll
    for (int k = 0; k < ACC; ++k) {
      double2 t = make_double2(0.0, 0.0);
      if (k < K) {
        if (a.q_prev) {
          CHECK16(a.q_prev + (size_t)k * a.pitch + c0, 2);
          t = *reinterpret_cast<const double2 *>(a.q_prev + (size_t)k * a.pitch + c0);
        } else if (k == 0) {
          t = make_double2(1.0, 1.0);
        }
      }
      if (!valid0) t.x = 0.0;
      if (!valid1) t.y = 0.0;
      CHECK16(my_q + (size_t)k * kTilePitch + 2 * lane, 3);
      *reinterpret_cast<double2 *>(my_q + (size_t)k * kTilePitch + 2 * lane) = t;
    }
  }

  const unsigned char *my_ring = smem_raw + 16 * threadIdx.x;  // this thread's 16-byte column of every ring row
  const double df = a.settle_df;
  const int num_groups = (a.G + kGroup - 1) / kGroup;
  const int grp_begin = blockIdx.x * a.subtiles_per_cta;
  const int grp_end = min(grp_begin + a.subtiles_per_cta, num_groups);
  const double *tile_src = a.v_next + tile * kTilePaths;  // column block of this CTA in every row

  // ring bookkeeping is uniform across the CTA; only thread 0 issues the copies
  int have_hi = a.win_lo[grp_begin] - 1;
  auto issue_rows = [&](int grp) {
    const int hi = a.win_hi[grp];
#ifdef LSMC_RING_NO_TMA
    for (int r = have_hi + 1; r <= hi; ++r)
      for (int i = threadIdx.x; i < kTilePaths; i += blockDim.x) s_ring[(size_t)(r % R) * kTilePaths + i] = tile_src[(size_t)r * a.pitch + i];
    if (false) {
#else
    if (threadIdx.x == 0) {
#endif
      const int count = hi - have_hi;
      // two barriers used alternately: group i completes on barrier (i - grp_begin) & 1, whose previous phase (group
      // i-2) has been waited for by every thread before this is issued
      uint64_t *bar = s_bar + ((grp - grp_begin) & 1);
      mbar_expect_tx(bar, (uint32_t)(count > 0 ? count : 0) * kRowBytes);
      for (int r = have_hi + 1; r <= hi; ++r)
        tma_load_row(s_ring + (size_t)(r % R) * kTilePaths, tile_src + (size_t)r * a.pitch, kRowBytes, bar);
    }
    if (hi > have_hi) have_hi = hi;
  };
#ifdef LSMC_RING_CHECK
  if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0)
    printf("ring kernel: R=%d D=%d K=%d G=%d pitch=%lld smem=%p bar=%p sub=%d win=[%d,%d]\n", R, D, K, a.G, (long long)a.pitch,
           (void *)smem_raw, (void *)s_bar, a.subtiles_per_cta, a.win_lo[0], a.win_hi[0]);
  if (((unsigned long long)s_bar) & 7ull) printf("misaligned mbarrier %p\n", (void *)s_bar);
#endif
  if (threadIdx.x == 0) {
    mbar_init(s_bar, 1);
    mbar_init(s_bar + 1, 1);
  }
  __syncthreads();
  issue_rows(grp_begin);

  for (int grp = grp_begin; grp < grp_end; ++grp) {
    const int g0 = grp * kGroup;
    const int gcount = min(kGroup, a.G - g0);
    __syncthreads();  // (A) every warp is done with the previous group's tables, tile and partial sums
    if (a.accumulate && grp > grp_begin) {
      const int gprev = (grp - 1) * kGroup;
      for (int i = threadIdx.x; i < kGroup * ACC; i += blockDim.x) {
        const int g = i / ACC, k = i - g * ACC;
        double sum = s_acc[((size_t)0 * ACC + k) * kGroup + g];
        for (int w = 1; w < kWarps; ++w) sum += s_acc[((size_t)w * ACC + k) * kGroup + g];
        a.partials[((size_t)tile * a.G + gprev + g) * ACC + k] = sum;  // gprev + g < G: previous groups are full
      }
    }
    for (int i = threadIdx.x; i < gcount * D; i += blockDim.x) {
      const int g = i / D;
      if (i - g * D < a.num_decisions[g0 + g]) {
        const DevDecision e = a.decisions[(size_t)g0 * D + i];
        CHECK16(s_a + i, 4);
        CHECK16(s_b + i, 5);
        s_a[i] = DecisionA{-e.volume, -e.consumed, e.cost_npv, e.lower_w};
        s_b[i] = DecisionR{e.upper_w, (uint32_t)(e.lower % R) * kRowBytes, (uint32_t)(e.upper % R) * kRowBytes};
        fold_coefficients<KC>(e, a.gamma, K, s_coef + (size_t)i * KC);
      }
    }
    for (int i = threadIdx.x; i < gcount; i += blockDim.x) {
      s_inv_cost[i] = a.inventory_cost[g0 + i];
      s_nd[i] = a.num_decisions[g0 + i];
    }
    if (grp + 1 < grp_end) issue_rows(grp + 1);  // only overwrites rows below this group's window
    __syncthreads();  // (B) tables visible
#ifndef LSMC_RING_NO_TMA
    mbar_wait(s_bar + ((grp - grp_begin) & 1), ((grp - grp_begin) >> 1) & 1);  // this group's rows have landed
#endif

    for (int gi = 0; gi < kGroup; ++gi) {
      double v0 = 0.0, v1 = 0.0;
      if (gi < gcount) {
        const int nd = s_nd[gi];
        const double inv_cost = s_inv_cost[gi];
        double best0 = 0.0, best1 = 0.0, adj0 = 0.0, adj1 = 0.0;
        if (D <= 3) {
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            if (d < nd) {
              const DecisionA da = s_a[gi * D + d];
              const DecisionR db = s_b[gi * D + d];
              const double2 lo = *reinterpret_cast<const double2 *>(my_ring + db.lower_off);
              const double2 up = *reinterpret_cast<const double2 *>(my_ring + db.upper_off);
              evaluate_decision<KP>(d == 0, da, db.upper_w, s_coef + (size_t)(gi * D + d) * KC, lo, up, q0, q1, price0, price1, df,
                                    inv_cost, best0, best1, adj0, adj1);
            }
          }
        } else {
          for (int d = 0; d < nd; ++d) {
            const DecisionA da = s_a[gi * D + d];
            const DecisionR db = s_b[gi * D + d];
            const double2 lo = *reinterpret_cast<const double2 *>(my_ring + db.lower_off);
            const double2 up = *reinterpret_cast<const double2 *>(my_ring + db.upper_off);
            evaluate_decision<KP>(d == 0, da, db.upper_w, s_coef + (size_t)(gi * D + d) * KC, lo, up, q0, q1, price0, price1, df,
                                  inv_cost, best0, best1, adj0, adj1);
          }
        }
        v0 = valid0 ? best0 + adj0 : 0.0;
        v1 = valid1 ? best1 + adj1 : 0.0;
        CHECK16(a.v_this + (size_t)(g0 + gi) * a.pitch + n0, 8);
        if (valid0) *reinterpret_cast<double2 *>(a.v_this + (size_t)(g0 + gi) * a.pitch + n0) = make_double2(v0, v1);
      }
      CHECK16(my_v + (size_t)gi * kTilePitch + 2 * lane, 9);
      if (a.accumulate) *reinterpret_cast<double2 *>(my_v + (size_t)gi * kTilePitch + 2 * lane) = make_double2(v0, v1);
    }

#ifdef LSMC_RING_NO_MMA
    if (false) {
#else
    if (a.accumulate) {
#endif
      __syncwarp();
      const int row = lane >> 2, col = lane & 3;
#pragma unroll
      for (int mb = 0; mb < ACC; mb += 8) {
        double acc0 = 0.0, acc1 = 0.0;
        const double *qa = my_q + (size_t)(mb + row) * kTilePitch + col;
        const double *vb = my_v + (size_t)row * kTilePitch + col;
#pragma unroll 4
        for (int p = 0; p < kWarpPaths; p += 4) dmma_m8n8k4(acc0, acc1, qa[p], vb[p]);
        double *dst = s_acc + ((size_t)warp * ACC + mb + row) * kGroup + 2 * col;
        dst[0] = acc0;
        dst[1] = acc1;
      }
      __syncwarp();
    }
  }

  if (a.accumulate && grp_end > grp_begin) {
    __syncthreads();
    const int gprev = (grp_end - 1) * kGroup;
    const int gcount = min(kGroup, a.G - gprev);
    for (int i = threadIdx.x; i < gcount * ACC; i += blockDim.x) {
      const int g = i / ACC, k = i - g * ACC;
      double sum = s_acc[((size_t)0 * ACC + k) * kGroup + g];
      for (int w = 1; w < kWarps; ++w) sum += s_acc[((size_t)w * ACC + k) * kGroup + g];
      a.partials[((size_t)tile * a.G + gprev + g) * ACC + k] = sum;
    }
  }
}

template <int KP>
void launch_ring(const BackwardStepArgs &a, dim3 grid, size_t smem, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(backward_ring_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    configured = true;
  }
  backward_ring_kernel<KP><<<grid, dim3(kThreads), smem, stream>>>(a);
}

}  // namespace

size_t backward_ring_smem_bytes(int KP, int D, int ring_rows) {
  const int acc = KP <= 8 ? 8 : 16, kc = (KP + 1) & ~1;
  return (size_t)ring_rows * kRowBytes + (size_t)kGroup * D * 64 +
         sizeof(double) * ((size_t)kGroup * D * kc + (size_t)kWarps * 8 * kTilePitch + (size_t)kWarps * acc * kTilePitch +
                           (size_t)kWarps * acc * kGroup + kGroup) +
         sizeof(int32_t) * kGroup + 16 + 128;
}

void launch_backward_ring(const BackwardStepArgs &a, int gsplit, int64_t num_tiles, cudaStream_t stream) {
  const int KP = padded_basis(a.K);
  const dim3 grid(gsplit, (unsigned)num_tiles);
  const size_t smem = backward_ring_smem_bytes(KP, a.D, a.ring_rows);
  switch (KP) {
    case 1: launch_ring<1>(a, grid, smem, stream); break;
    case 2: launch_ring<2>(a, grid, smem, stream); break;
    case 3: launch_ring<3>(a, grid, smem, stream); break;
    case 4: launch_ring<4>(a, grid, smem, stream); break;
    case 5: launch_ring<5>(a, grid, smem, stream); break;
    case 6: launch_ring<6>(a, grid, smem, stream); break;
    case 7: launch_ring<7>(a, grid, smem, stream); break;
    case 8: launch_ring<8>(a, grid, smem, stream); break;
    case 12: launch_ring<12>(a, grid, smem, stream); break;
    default: launch_ring<16>(a, grid, smem, stream); break;
  }
}

}  // namespace lsmc
