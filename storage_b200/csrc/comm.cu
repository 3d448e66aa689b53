// comm.cu — see comm.hpp.
#include "comm.hpp"

#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "host_model.hpp"

namespace lsmc {

namespace {

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi &nccl() {
  static NcclApi api;
  if (api.handle) return api;
  // inside a process that imported torch the bundled libnccl.so.2 is already mapped and is found by soname
  const char *override_path = std::getenv("LSMC_NCCL_LIBRARY");
  const char *candidates[] = {override_path, "libnccl.so.2", "libnccl.so"};
  for (const char *c : candidates) {
    if (!c) continue;
    api.handle = dlopen(c, RTLD_NOW | RTLD_GLOBAL);
    if (api.handle) break;
  }
  if (!api.handle)
    throw Error(LSMC_ERR_CUDA, std::string("cannot load NCCL (set LSMC_NCCL_LIBRARY to the path of libnccl.so.2): ") + dlerror());
  auto sym = [&](const char *name) {
    void *p = dlsym(api.handle, name);
    if (!p) throw Error(LSMC_ERR_CUDA, std::string("NCCL symbol missing: ") + name);
    return p;
  };
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
  api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  return api;
}

void nccl_ok(ncclResult_t r, const char *what) {
  if (r != ncclSuccess) throw Error(LSMC_ERR_CUDA, std::string(what) + " failed: " + nccl().GetErrorString(r));
}

static_assert(sizeof(ncclUniqueId) <= LSMC_COMM_ID_BYTES, "LSMC_COMM_ID_BYTES too small for ncclUniqueId");

// Communicators are expensive to create (~1 s) and a host typically runs many valuations on the same set of ranks: a
// communicator is created once per unique id and shared by every handle that joins with that id (kept until exit).
std::mutex g_comm_mutex;
std::map<std::string, ncclComm_t> g_comms;

}  // namespace

Comm::~Comm() {
  release_peer_exchange();
  if (d_send_) cudaFree(d_send_);
  if (d_recv_) cudaFree(d_recv_);
}

void Comm::release_peer_exchange() {
  for (size_t r = 0; r < peer_base_.size(); ++r)
    if ((int)r != rank_ && peer_base_[r]) cudaIpcCloseMemHandle(peer_base_[r]);
  peer_base_.clear();
  if (d_peer_base_) cudaFree(d_peer_base_);
  if (xchg_local_) cudaFree(xchg_local_);
  d_peer_base_ = nullptr;
  xchg_local_ = nullptr;
}

bool Comm::setup_peer_exchange(size_t capacity_doubles, cudaStream_t stream) {
  if (world_ == 1) return false;
  if (const char *env = std::getenv("LSMC_PEER_EXCHANGE"))  // test hook: 0 keeps the NCCL all-gather path
    if (env[0] == '0') return false;
  release_peer_exchange();
  xchg_capacity_ = capacity_doubles;
  const size_t bytes = xchg_buffer_bytes(world_, capacity_doubles);
  static_assert(sizeof(cudaIpcMemHandle_t) % sizeof(double) == 0, "IPC handles travel as doubles through the all-gather");
  constexpr size_t kHandleDoubles = sizeof(cudaIpcMemHandle_t) / sizeof(double);
  // every rank takes part in the handle all-gather even when its own allocation failed (an all-zero handle says so):
  // the decision to use the peer path must be the same on every rank
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof(mine));
  bool ok = cudaMalloc(&xchg_local_, bytes) == cudaSuccess && cudaMemset(xchg_local_, 0, bytes) == cudaSuccess &&
            cudaDeviceSynchronize() == cudaSuccess && cudaIpcGetMemHandle(&mine, xchg_local_) == cudaSuccess;
  if (!ok) {
    cudaGetLastError();
    std::memset(&mine, 0, sizeof(mine));
  }
  std::vector<double> send(kHandleDoubles + 1), all;
  std::memcpy(send.data(), &mine, sizeof(mine));
  send[kHandleDoubles] = ok ? 1.0 : 0.0;
  all_gather_host(send.data(), send.size(), all, stream);
  for (int r = 0; r < world_; ++r) ok = ok && all[(size_t)r * send.size() + kHandleDoubles] == 1.0;
  // second round: did every rank manage to map every peer?
  std::vector<char *> bases(world_, nullptr);
  bool mapped = ok;
  if (ok) {
    for (int r = 0; r < world_ && mapped; ++r) {
      if (r == rank_) {
        bases[r] = xchg_local_;
        continue;
      }
      cudaIpcMemHandle_t h;
      std::memcpy(&h, &all[(size_t)r * send.size()], sizeof(h));
      void *p = nullptr;
      if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        mapped = false;
      }
      bases[r] = static_cast<char *>(p);
    }
  }
  double flag = mapped ? 1.0 : 0.0;
  std::vector<double> flags;
  all_gather_host(&flag, 1, flags, stream);
  bool all_mapped = true;
  for (double f : flags) all_mapped = all_mapped && f == 1.0;
  if (all_mapped && cudaMalloc(&d_peer_base_, world_ * sizeof(char *)) == cudaSuccess &&
      cudaMemcpy(d_peer_base_, bases.data(), world_ * sizeof(char *), cudaMemcpyHostToDevice) == cudaSuccess) {
    peer_base_ = bases;
    xchg_seq_ = 0;
    return true;
  }
  for (int r = 0; r < world_; ++r)
    if (r != rank_ && bases[r]) cudaIpcCloseMemHandle(bases[r]);
  if (xchg_local_) cudaFree(xchg_local_);
  xchg_local_ = nullptr;
  return false;
}

XchgArgs Comm::next_exchange(int32_t *error_flag) {
  XchgArgs x{};
  x.peers = d_peer_base_;
  x.world = world_;
  x.rank = rank_;
  x.capacity = xchg_capacity_;
  x.seq = ++xchg_seq_;
  static const unsigned long long timeout_ns = [] {
    const char *env = std::getenv("LSMC_PEER_TIMEOUT_MS");
    return (unsigned long long)(env ? std::atof(env) : 10000.0) * 1000000ull;
  }();
  x.timeout_ns = timeout_ns;
  x.error_flag = error_flag;
  return x;
}

void Comm::unique_id(void *id_out) {
  if (!id_out) throw Error(LSMC_ERR_ARGUMENT, "id_out is null");
  ncclUniqueId id;
  nccl_ok(nccl().GetUniqueId(&id), "ncclGetUniqueId");
  std::memset(id_out, 0, LSMC_COMM_ID_BYTES);
  std::memcpy(id_out, &id, sizeof(id));
}

void Comm::init(const void *id, int rank, int world, int64_t total_sims) {
  if (world < 1 || rank < 0 || rank >= world) throw Error(LSMC_ERR_ARGUMENT, "invalid rank / world size");
  if (nccl_comm_) throw Error(LSMC_ERR_INVALID_OPERATION, "communicator already initialised for this handle");
  world_ = world;
  rank_ = rank;
  total_sims_ = total_sims;
  if (world == 1) return;
  if (!id) throw Error(LSMC_ERR_ARGUMENT, "communicator id is null");
  ncclUniqueId uid;
  std::memcpy(&uid, id, sizeof(uid));
  const std::string key(reinterpret_cast<const char *>(&uid), sizeof(uid));
  std::lock_guard<std::mutex> lock(g_comm_mutex);
  auto it = g_comms.find(key);
  if (it == g_comms.end()) {
    ncclComm_t comm;
    nccl_ok(nccl().CommInitRank(&comm, world, uid, rank), "ncclCommInitRank");
    it = g_comms.emplace(key, comm).first;
  }
  nccl_comm_ = it->second;
}

void Comm::ensure_staging(size_t count) {
  if (count <= staging_) return;
  if (d_send_) cudaFree(d_send_);
  if (d_recv_) cudaFree(d_recv_);
  if (cudaMalloc(&d_send_, count * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&d_recv_, count * world_ * sizeof(double)) != cudaSuccess)
    throw Error(LSMC_ERR_CUDA, "cudaMalloc of communication staging failed");
  staging_ = count;
}

void Comm::all_gather(const double *send, double *recv, size_t count, cudaStream_t stream) {
  if (world_ == 1) {
    if (send != recv) cudaMemcpyAsync(recv, send, count * sizeof(double), cudaMemcpyDeviceToDevice, stream);
    return;
  }
  nccl_ok(nccl().AllGather(send, recv, count, ncclFloat64, static_cast<ncclComm_t>(nccl_comm_), stream), "ncclAllGather");
}

void Comm::all_gather_host(const double *values, size_t n, std::vector<double> &all, cudaStream_t stream) {
  all.assign(n * world_, 0.0);
  if (world_ == 1) {
    std::copy(values, values + n, all.begin());
    return;
  }
  if (n == 0) return;
  ensure_staging(n);
  auto ok = [](cudaError_t e, const char *what) {
    if (e != cudaSuccess) throw Error(LSMC_ERR_CUDA, std::string(what) + " failed in all_gather_host: " + cudaGetErrorString(e));
  };
  ok(cudaMemcpyAsync(d_send_, values, n * sizeof(double), cudaMemcpyHostToDevice, stream), "cudaMemcpyAsync");
  all_gather(d_send_, d_recv_, n, stream);
  ok(cudaMemcpyAsync(all.data(), d_recv_, all.size() * sizeof(double), cudaMemcpyDeviceToHost, stream), "cudaMemcpyAsync");
  ok(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
}

void Comm::all_reduce_host(double *values, size_t n, cudaStream_t stream) {
  if (world_ == 1 || n == 0) return;
  std::vector<double> all;
  all_gather_host(values, n, all, stream);
  for (size_t i = 0; i < n; ++i) {
    double t = 0.0;
    for (int r = 0; r < world_; ++r) t += all[(size_t)r * n + i];  // rank order: identical on every rank
    values[i] = t;
  }
}

}  // namespace lsmc
