// kernels.hpp — launch interface between lsmc_api.cu (host orchestration) and kernels.cu.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>

#include "device_types.hpp"

namespace lsmc {

struct GramArgs {
  int64_t N;
  int64_t chunk_paths;     // paths per CTA
  int K, F;
  const BasisSpec *specs;  // [num_slots] device array: one per regression period
  const int32_t *powers;   // [K][1+F]
  double *partials;        // [num_slots][num_chunks][KP*(KP+1)/2][2]
};

struct CholeskyArgs {
  int K;
  int num_ranks, num_slots, num_chunks;
  const double *partials;  // [num_ranks][num_slots][num_chunks][NP][2], summed in (rank, chunk) order
  double *rinv;            // [num_slots][K*K*2]
  int32_t *error_flag;
};

struct BasisArgs {
  int64_t N, pitch;
  int K, F;
  BasisSpec spec;          // regression simulations of the period + its R^-1
  const int32_t *powers;
  double *q_out;           // [K][pitch]
};

struct ReduceArgs {
  const double *partials;  // [G][num_chunks][acc_k]: per path tile sums, written by the backward kernels
  int num_chunks, G, K, acc_k;
  int mean_mode;
  int64_t total_paths;
  double *out;             // [G][K]
};

// Inputs of fold_tables_kernel (backward_sweep.cu): one record per grid point of the step with everything that does not
// depend on the path.
struct FoldArgs {
  int G, Gn, K;  // grid points of this period and of the next, basis functions
  int64_t pitch;
  const DevDecision *decisions;  // [G][3]
  const int32_t *num_decisions;  // [G]
  const double *inventory_cost;  // [G]
  const double *gamma;           // [Gn][K]
  double settle_df;
  int fused_payoff;
  int ring_rows;                 // > 0: row offsets are written as slots of a ring of this many rows (row-ring sweep kernel)
  int format;                    // 0: one PointTable per grid point (backward_sweep.cu); 1: one MmaBlockRecord per block of 8 grid
                                 //    points (backward_sweep_mma.cu, mma_records.cuh)
  int row_pitch_bytes;           // format 1: distance between two rows of the shared-memory ring
  void *out;                     // format 0: [G] records of point_table_bytes(K) each; format 1: [ceil(G/8)] block records
};

// Peer-memory exchange of the per-step regression sums between the ranks of a sharded valuation (comm.cu PeerExchange).
// Every rank owns one exchange buffer and maps every other rank's (CUDA IPC over NVLink):
//   [world] arrival flags, 128 bytes apart: flags[s] = sequence number of the last exchange rank s has fully written here
//   [2][world][capacity] doubles: inbox, one slot per (parity of the sequence number, source rank)
struct XchgArgs {
  char *const *peers;     // [world] base of every rank's exchange buffer as mapped here (own buffer at [rank])
  int world, rank;
  size_t capacity;        // doubles per inbox slot
  unsigned long long seq; // this exchange's sequence number (same on every rank, increasing)
  unsigned long long timeout_ns;  // give up waiting for a peer after this long (flags an error instead of hanging the GPU)
  int32_t *error_flag;
};
constexpr size_t kXchgFlagStride = 128;
__host__ __device__ inline size_t xchg_data_offset(int world) { return ((size_t)world * kXchgFlagStride + 255) / 256 * 256; }
__host__ __device__ inline size_t xchg_buffer_bytes(int world, size_t capacity) { return xchg_data_offset(world) + 2 * (size_t)world * capacity * sizeof(double); }

// Raises a kernel's dynamic shared-memory limit, once per device and size (the attribute is sticky; setting it on every
// launch costs host time in the per-step loop), and reports a refusal instead of letting the launch fail later.
template <auto Kernel>
inline void ensure_dynamic_smem(size_t smem) {
  if (smem + 4096 <= 48 * 1024) return;  // the default limit counts the kernel's static shared memory too: leave room for it
  constexpr int kDevices = 64;
  static std::atomic<size_t> granted[kDevices] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  const bool tracked = dev >= 0 && dev < kDevices;
  if (tracked && granted[dev].load(std::memory_order_relaxed) >= smem) return;
  const cudaError_t e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess)
    throw std::runtime_error(std::string("cudaFuncSetAttribute(MaxDynamicSharedMemorySize = ") + std::to_string(smem) + ") failed: " + cudaGetErrorString(e));
  if (tracked) granted[dev].store(smem, std::memory_order_relaxed);
}

int padded_basis(int K);
inline int acc_width(int K) { return padded_basis(K) <= 8 ? 8 : 16; }
size_t backward_smem_bytes(int KP, int D);
// One period's tables of the forward kernel in shared memory (byte offsets inside a table buffer, every part 16-byte aligned)
struct ForwardTableLayout {
  uint32_t params, gamma, grid, rinv, first, bytes;
};
__host__ __device__ inline ForwardTableLayout forward_table_layout(int K, int Gmax, int first_pitch) {
  ForwardTableLayout l;
  l.params = 0;
  l.gamma = (uint32_t)sizeof(DevPeriodParams);
  l.grid = l.gamma + (((uint32_t)Gmax * K * 8u + 15u) & ~15u);
  l.rinv = l.grid + (((uint32_t)Gmax * 8u + 15u) & ~15u);
  l.first = l.rinv + (uint32_t)K * K * 16u;
  l.bytes = l.first + (uint32_t)first_pitch;
  return l;
}
inline int forward_first_pitch(int Gmax) { return (Gmax + 15) & ~15; }  // row pitch of ForwardArgs::bisect_first: whole 16-byte copies
size_t forward_smem_bytes(int K, int F, int Gmax, int first_pitch, int table_buffers, int sums_width, int tiles_per_cta);

void launch_gram(const GramArgs &a, int num_chunks, int num_slots, cudaStream_t stream);
void launch_cholesky(const CholeskyArgs &a, int num_slots, cudaStream_t stream);
void launch_backward_step(const BackwardStepArgs &a, int gsplit, int64_t num_tiles, cudaStream_t stream);
void launch_basis(const BasisArgs &a, cudaStream_t stream);
// Three-decision sweep kernel (backward_sweep.cu): gsplit CTAs per 128-path tile, each sweeping a.subtiles_per_cta blocks of 8
// grid points; partial sums are laid out [tile of 128 paths][G][acc].
size_t point_table_bytes(int K);
size_t backward_sweep_smem_bytes(int K);
size_t backward_sweep_ring_smem_bytes(int K, int ring_rows, int blocks_per_cta);
void launch_fold_tables(const FoldArgs &f, cudaStream_t stream);
// reduce_partials + fold_tables in one launch (single GPU): the last CTA to finish folds; `counter` is a zeroed device word
void launch_reduce_fold(const ReduceArgs &r, const FoldArgs &f, unsigned int *counter, cudaStream_t stream);
// sharded runs: reduce_partials + exchange with the other ranks over peer memory + rank-ordered sum + fold_tables in one launch
void launch_exchange_fold(const ReduceArgs &r, const FoldArgs &f, const XchgArgs &x, unsigned int *counter, cudaStream_t stream);
void launch_backward_sweep(const BackwardStepArgs &a, int gsplit, int64_t num_tiles, cudaStream_t stream);
// FP64-mma sweep kernel (backward_sweep_mma.cu): fused-payoff form only; look-ahead of kMmaLookAhead blocks in the row ring
#ifndef LSMC_MMA_LOOKAHEAD
#define LSMC_MMA_LOOKAHEAD 2
#endif
constexpr int kMmaLookAhead = LSMC_MMA_LOOKAHEAD;  // generations of rows requested ahead of the block being evaluated
#ifndef LSMC_MMA_SLACK
#define LSMC_MMA_SLACK 1
#endif
// Generations kept beyond the look-ahead: the slot generation b + look-ahead goes into was last read by block b - 1 - slack,
// so with slack the issuing warp does not have to wait for the slowest warp of the block just before its own.
constexpr int kMmaSlack = LSMC_MMA_SLACK;
constexpr int kMmaRowPitchBytes = kSweepTilePaths * 8 + 64;  // consecutive ring rows start in opposite halves of the 32 banks
size_t mma_block_record_bytes(int K);
size_t backward_sweep_mma_smem_bytes(int K, int ring_rows, int blocks_per_cta);
// a.num_tiles, a.whole_tiles and a.tail_split describe the grid (device_types.hpp)
void launch_backward_sweep_mma(const BackwardStepArgs &a, cudaStream_t stream);
int backward_sweep_mma_ctas_per_sm();
void launch_reduce_partials(const ReduceArgs &a, cudaStream_t stream);
void launch_forward(const ForwardArgs &a, int num_ctas, cudaStream_t stream);
// out[i] = sum over ranks r of all[r][i] (rank order); mean_mode: component 0 of every K-vector divided by total_paths, rest 0
void launch_combine_ranks(const double *all, int world, int count, int K, int mean_mode, double total_paths, double *out,
                          cudaStream_t stream);
void launch_reduce_step_sums(const double *in, int P, int num_ctas, int width, double *out, cudaStream_t stream);
void launch_sum(const double *x, int64_t n, double shift, int power, double *out, int blocks, cudaStream_t stream);

}  // namespace lsmc
