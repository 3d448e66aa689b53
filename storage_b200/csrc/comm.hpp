// comm.hpp — multi-GPU plumbing: one process per GPU, paths sharded across ranks.  The only exchanges on
// the path are (1) the double-double Gram partials of the pre-pass, (2) one gather of G'xK regression sums per backward
// step and (3) the final per-period / NPV sums.  All of them are all-gathers followed by a sum in rank order, so every
// rank computes bit-identical regression coefficients.  NCCL is loaded with dlopen the first time a communicator is
// requested: the library itself has no link-time dependency on it and single-GPU runs never touch it.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <vector>

#include "kernels.hpp"

namespace lsmc {

class Comm {
 public:
  Comm() = default;
  ~Comm();
  Comm(const Comm &) = delete;
  Comm &operator=(const Comm &) = delete;

  static void unique_id(void *id_out);  // rank 0: LSMC_COMM_ID_BYTES bytes to broadcast to the other ranks
  void init(const void *id, int rank, int world, int64_t total_sims);

  int world() const { return world_; }
  int rank() const { return rank_; }
  int64_t total_sims(int64_t local) const { return world_ > 1 ? total_sims_ : local; }

  // recv[r][count] <- rank r's send[count] (device pointers), on `stream`
  void all_gather(const double *send, double *recv, size_t count, cudaStream_t stream);
  // sums n host doubles over ranks in rank order (no-op for a single rank); synchronises `stream`
  void all_reduce_host(double *values, size_t n, cudaStream_t stream);
  // all[r][n] <- rank r's values[n] (host); synchronises `stream`
  void all_gather_host(const double *values, size_t n, std::vector<double> &all, cudaStream_t stream);

  // Peer-memory exchange of the per-step regression sums (kernels.hpp XchgArgs): every rank allocates one buffer, the
  // ranks swap CUDA IPC handles through the communicator and map each other's buffers.  Returns false (and leaves the NCCL
  // all-gather path in use) when the buffers cannot be mapped, e.g. ranks on different nodes or IPC disabled in the container.
  bool setup_peer_exchange(size_t capacity_doubles, cudaStream_t stream);
  bool peer_exchange_ready() const { return !peer_base_.empty(); }
  XchgArgs next_exchange(int32_t *error_flag);  // arguments of the next exchange; every rank must call it equally often

 private:
  int world_ = 1, rank_ = 0;
  int64_t total_sims_ = 0;
  void *nccl_comm_ = nullptr;
  double *d_send_ = nullptr, *d_recv_ = nullptr;
  size_t staging_ = 0;
  void ensure_staging(size_t count);
  // peer exchange
  char *xchg_local_ = nullptr;
  std::vector<char *> peer_base_;  // [world] mapped bases (own buffer at [rank])
  char **d_peer_base_ = nullptr;
  size_t xchg_capacity_ = 0;
  unsigned long long xchg_seq_ = 0;
  void release_peer_exchange();
};

}  // namespace lsmc
