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
  if (d_send_) cudaFree(d_send_);
  if (d_recv_) cudaFree(d_recv_);
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

void Comm::all_reduce_host(double *values, size_t n, cudaStream_t stream) {
  if (world_ == 1 || n == 0) return;
  ensure_staging(n);
  cudaMemcpyAsync(d_send_, values, n * sizeof(double), cudaMemcpyHostToDevice, stream);
  all_gather(d_send_, d_recv_, n, stream);
  std::vector<double> all(n * world_);
  cudaMemcpyAsync(all.data(), d_recv_, all.size() * sizeof(double), cudaMemcpyDeviceToHost, stream);
  if (cudaStreamSynchronize(stream) != cudaSuccess) throw Error(LSMC_ERR_CUDA, "stream synchronisation failed in all_reduce_host");
  for (size_t i = 0; i < n; ++i) {
    double t = 0.0;
    for (int r = 0; r < world_; ++r) t += all[(size_t)r * n + i];  // rank order: identical on every rank
    values[i] = t;
  }
}

}  // namespace lsmc
