// lsmc_api.cu — the C ABI of include/lsmc_b200.h: handle life cycle, host orchestration of the backward induction
// and forward simulation on one B200, result marshalling.  No CPU fallback: every compute entry point fails with
// LSMC_ERR_CUDA when there is no usable device.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <climits>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <exception>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "comm.hpp"
#include "device_types.hpp"
#include "host_model.hpp"
#include "kernels.hpp"
#include "simulate.hpp"

using namespace lsmc;

namespace {

thread_local std::string g_create_error;

struct CudaFailure : std::runtime_error {
  explicit CudaFailure(const std::string &m) : std::runtime_error(m) {}
};

#define CUDA_OK(expr)                                                                                          \
  do {                                                                                                         \
    cudaError_t err__ = (expr);                                                                                \
    if (err__ != cudaSuccess)                                                                                  \
      throw CudaFailure(std::string(#expr) + " failed: " + cudaGetErrorString(err__) + " (" __FILE__ ":" +     \
                        std::to_string(__LINE__) + ")");                                                       \
  } while (0)

template <class T>
struct DeviceBuffer {
  T *ptr = nullptr;
  size_t count = 0;
  DeviceBuffer() = default;
  DeviceBuffer(const DeviceBuffer &) = delete;
  DeviceBuffer &operator=(const DeviceBuffer &) = delete;
  ~DeviceBuffer() { release(); }
  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    count = 0;
  }
  void allocate(size_t n) {  // contents are unspecified afterwards; an existing buffer of the right size is kept (handle re-use)
    if (n == 0) n = 1;
    if (ptr && count == n) return;
    release();
    CUDA_OK(cudaMalloc(&ptr, n * sizeof(T)));
    count = n;
  }
  void upload(const std::vector<T> &host) {
    allocate(host.size());
    if (!host.empty()) CUDA_OK(cudaMemcpy(ptr, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice));
  }
};

// Host <-> device movement of whole simulation / result panels (the caller-supplied-simulation overload at production size is
// 58 GB per path set).  The caller's arrays are pageable, so they are staged through two pinned buffers: while
// the DMA engine moves chunk i, a few host threads copy chunk i+1 between the caller's array and the other buffer.
class PinnedStager {
 public:
  static constexpr size_t kChunkBytes = size_t(64) << 20;
  ~PinnedStager() {
    for (int i = 0; i < 2; ++i) {
      if (buf_[i]) cudaFreeHost(buf_[i]);
      if (ev_[i]) cudaEventDestroy(ev_[i]);
    }
  }
  void to_device(void *dst, const void *src, size_t bytes, cudaStream_t st) {
    ensure();
    char *d = static_cast<char *>(dst);
    const char *s = static_cast<const char *>(src);
    int i = 0;
    for (size_t off = 0; off < bytes; off += kChunkBytes, i ^= 1) {
      const size_t len = std::min(kChunkBytes, bytes - off);
      CUDA_OK(cudaEventSynchronize(ev_[i]));  // the copy that last used this buffer has left it
      host_copy(buf_[i], s + off, len);
      CUDA_OK(cudaMemcpyAsync(d + off, buf_[i], len, cudaMemcpyHostToDevice, st));
      CUDA_OK(cudaEventRecord(ev_[i], st));
    }
    CUDA_OK(cudaStreamSynchronize(st));
  }
  void to_host(void *dst, const void *src, size_t bytes, cudaStream_t st) {
    ensure();
    char *d = static_cast<char *>(dst);
    const char *s = static_cast<const char *>(src);
    size_t pending_off = 0, pending_len = 0;
    int i = 0, pending_buf = -1;
    for (size_t off = 0; off < bytes; off += kChunkBytes, i ^= 1) {
      const size_t len = std::min(kChunkBytes, bytes - off);
      CUDA_OK(cudaMemcpyAsync(buf_[i], s + off, len, cudaMemcpyDeviceToHost, st));
      CUDA_OK(cudaEventRecord(ev_[i], st));
      if (pending_buf >= 0) {  // drain the previous chunk while this one is in flight
        CUDA_OK(cudaEventSynchronize(ev_[pending_buf]));
        host_copy(d + pending_off, buf_[pending_buf], pending_len);
      }
      pending_buf = i, pending_off = off, pending_len = len;
    }
    if (pending_buf >= 0) {
      CUDA_OK(cudaEventSynchronize(ev_[pending_buf]));
      host_copy(d + pending_off, buf_[pending_buf], pending_len);
    }
  }

 private:
  void ensure() {
    for (int i = 0; i < 2; ++i) {
      if (!buf_[i]) CUDA_OK(cudaHostAlloc(&buf_[i], kChunkBytes, cudaHostAllocDefault));
      if (!ev_[i]) {
        CUDA_OK(cudaEventCreateWithFlags(&ev_[i], cudaEventDisableTiming));
      }
    }
  }
  static void host_copy(void *dst, const void *src, size_t bytes) {
    constexpr size_t kSlice = size_t(8) << 20;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t parts = std::min<size_t>({(bytes + kSlice - 1) / kSlice, size_t(hw), size_t(8)});
    if (parts <= 1) {
      std::memcpy(dst, src, bytes);
      return;
    }
    std::vector<std::thread> pool;
    const size_t per = (bytes / parts + 63) & ~size_t(63);
    for (size_t t = 1; t < parts; ++t) {
      const size_t off = t * per;
      if (off >= bytes) break;
      const size_t len = std::min(per, bytes - off);
      pool.emplace_back([=] { std::memcpy(static_cast<char *>(dst) + off, static_cast<const char *>(src) + off, len); });
    }
    std::memcpy(dst, src, std::min(per, bytes));
    for (auto &t : pool) t.join();
  }
  void *buf_[2] = {nullptr, nullptr};
  cudaEvent_t ev_[2] = {nullptr, nullptr};
};

struct SimSet {
  bool ready = false;
  const double *spot = nullptr;  // [rows][N] device
  const double *factors[LSMC_MAX_FACTORS] = {};
  DeviceBuffer<double> owned_spot;
  DeviceBuffer<double> owned_factors[LSMC_MAX_FACTORS];
  double ms_simulation = 0.0;
  double ms_upload = 0.0;  // host wall time of staging the caller's panels to the device
  // lsmc_set_sims_async: the panels travel on a worker thread with its own stream and pinned staging while the caller goes on
  // (the valuation simulations are not needed before the forward pass: their upload overlaps the backward induction)
  std::thread uploader;
  std::exception_ptr upload_error;
  std::unique_ptr<PinnedStager> upload_stager;
  cudaStream_t upload_stream = nullptr;
  void wait_upload() {  // joins the worker; rethrows what it threw
    if (uploader.joinable()) uploader.join();
    if (upload_error) {
      std::exception_ptr e = upload_error;
      upload_error = nullptr;
      std::rethrow_exception(e);
    }
  }
  ~SimSet() {
    if (uploader.joinable()) uploader.join();
    if (upload_stream) cudaStreamDestroy(upload_stream);
  }
};

}  // namespace

struct lsmc_handle {
  std::string error;
  int device = 0;
  // problem
  int64_t N = 0;
  int P = 0, K = 0, F = 0, KP = 0;
  bool current_active = false, discount_deltas = false, compat_terminal = true;
  uint32_t flags = 0;
  bool want_decisions = true;
  bool use_mma = false;  // three-decision tables in the fused-payoff form run on the FP64-mma sweep kernel
  double inventory = 0, tolerance = 1e-10, grid_spacing = 0;
  int extra = 0;
  std::vector<double> forward_price, settle_df;
  std::vector<int32_t> powers;
  std::unique_ptr<StorageModel> model;
  ValuationTables tables;
  int first_sim = 0, sim_rows = 0;
  int64_t pitch = 0;
  int Gmax = 0, D = 0, acc_k = 8;

  cudaStream_t stream = nullptr;
  PinnedStager stager;                         // pinned double buffer for whole-panel host <-> device copies
  cudaStream_t aux_stream = nullptr;          // basis rows of period p-3 are materialised here while step p-1 runs
  cudaEvent_t ev_basis[3] = {}, ev_step[3] = {}, ev_prepass = nullptr;
  std::vector<cudaEvent_t> step_events;  // two timing events per backward step, created once and kept for later runs
  cudaEvent_t ev_window[3] = {};        // completion of a window of backward steps: progress reports and cancellation follow the GPU
  SimSet sims[2];
  DeviceBuffer<double> v_a, v_b, grids, gamma, rinv, partials, gram_partials, inv_cost, pv, final_inv, step_sums, step_totals,
      block_sums, gather_buf, qbuf[3], ratchet_inv, ratchet_min, ratchet_max, ratchet_min_slope, ratchet_max_slope, poly_min, poly_max;
  DeviceBuffer<DevDecision> decisions;
  DeviceBuffer<double> point_tables;  // one step's per-grid-point records for the sweep kernel (kernels.hpp FoldArgs)
  DeviceBuffer<int32_t> num_decisions, d_powers, error_flag;
  DeviceBuffer<uint8_t> bisect_first;  // ForwardArgs::bisect_first
  DeviceBuffer<unsigned int> fold_counter;  // device-wide arrival counter of reduce_fold_kernel (self-resetting)
  std::vector<size_t> window_offset;  // per period: first block in sweep_blocks
  std::vector<int> sweep_ring_rows;   // per period: ring size of the row-ring sweep kernel (0: direct-load sweep kernel)
  DeviceBuffer<SweepBlock> sweep_blocks;  // [window_offset[p] + block] row windows of the row-ring sweep kernel
  DeviceBuffer<DevPeriodParams> period_params;
  DeviceBuffer<BasisSpec> gram_specs;
  DeviceBuffer<double> panels[7];  // inventory, volume, consumed, loss, net volume, pv, decision index
  // LSMC_STREAM_PANELS: the panels above are rings of 2 x stream_block periods; the forward simulation runs block by block and
  // every finished block is copied into the caller's host arrays (panel sinks) while the next one is computed
  bool stream_panels = false;
  int stream_block = 0;
  double *sinks[7] = {};
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_block[2] = {}, ev_copied[2] = {};
  PinnedStager sink_stager;
  std::vector<size_t> table_offset;  // per period: first grid point in the concatenated tables

  Comm comm;  // single rank unless lsmc_comm_init was called

  // results of the last run
  bool has_results = false;
  lsmc_results scalars{};
  std::vector<double> deltas, profile, pv_by_sim;
  std::vector<lsmc_trigger_prices> trigger;
  std::vector<double> h_gamma, h_rinv;
};

namespace {

int status_of(lsmc_handle *h, const std::exception &e) {
  if (auto *le = dynamic_cast<const Error *>(&e)) {
    if (h) h->error = le->what();
    else g_create_error = le->what();
    return le->status;
  }
  if (h) h->error = e.what();
  else g_create_error = e.what();
  return dynamic_cast<const CudaFailure *>(&e) ? LSMC_ERR_CUDA : LSMC_ERR_ARGUMENT;
}

void require_device(int device) {
  int count = 0;
  cudaError_t err = cudaGetDeviceCount(&count);
  if (err != cudaSuccess || count == 0)
    throw CudaFailure(std::string("no CUDA device available (") + cudaGetErrorString(err) +
                      "); liblsmc_b200 has no CPU fallback");
  if (device < 0 || device >= count) throw CudaFailure("CUDA device index out of range");
  CUDA_OK(cudaSetDevice(device));
}

BasisSpec basis_spec(const lsmc_handle &h, int which, int period) {
  BasisSpec b{};
  if (period == 0 && h.current_active) {
    b.mean_mode = 1;
    return b;
  }
  const SimSet &s = h.sims[which];
  const size_t row = (size_t)(period - h.first_sim) * h.N;
  b.spot = s.spot + row;
  for (int i = 0; i < h.F; ++i) b.factors[i] = s.factors[i] + row;
  b.rinv = h.rinv.ptr + (size_t)period * h.K * h.K * 2;
  b.mean_mode = 0;
  return b;
}

struct Timer {
  cudaEvent_t a, b;
  cudaStream_t s;
  explicit Timer(cudaStream_t stream) : s(stream) {
    cudaEventCreate(&a);
    cudaEventCreate(&b);
  }
  ~Timer() {
    cudaEventDestroy(a);
    cudaEventDestroy(b);
  }
  void start() { cudaEventRecord(a, s); }
  void stop() { cudaEventRecord(b, s); }
  double ms() {
    cudaEventSynchronize(b);
    float m = 0;
    cudaEventElapsedTime(&m, a, b);
    return m;
  }
};

double device_sum(lsmc_handle *h, const double *x, int64_t n, double shift, int power) {
  const int blocks = (int)std::min<int64_t>(512, std::max<int64_t>(1, (n + 4095) / 4096));
  if (h->block_sums.count < 512) h->block_sums.allocate(512);
  launch_sum(x, n, shift, power, h->block_sums.ptr, blocks, h->stream);
  h->scalars.kernel_launches++;
  std::vector<double> host(blocks);
  CUDA_OK(cudaMemcpyAsync(host.data(), h->block_sums.ptr, blocks * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CUDA_OK(cudaStreamSynchronize(h->stream));
  double t = 0.0;
  for (double v : host) t += v;
  return t;
}

// AverageContinuationValue (LSV.cs:785-805) from the per-period basis column means: mean_n(q_n . gamma) = qbar . gamma
double average_continuation(const lsmc_handle &h, int p, const double *qbar, double after) {
  const std::vector<double> &grid = h.tables.periods[p + 1].grid;
  int lo, hi;
  bisect_inventory_space(grid.data(), (int)grid.size(), after, h.tolerance, &lo, &hi);
  auto dot = [&](int g) {
    const double *gm = &h.h_gamma[((size_t)p * h.Gmax + g) * h.K];
    double s = 0.0;
    for (int k = 0; k < h.K; ++k) s += qbar[k] * gm[k];
    return s;
  };
  if (lo == hi) return dot(lo);
  const double lower_w = (grid[hi] - after) / (grid[hi] - grid[lo]);
  const double upper_w = 1.0 - lower_w;
  return dot(lo) * lower_w + dot(hi) * upper_w;
}

// Trigger prices of period p at the expected inventory (LSV.cs:523-592, 704-755)
void trigger_prices_for_period(const lsmc_handle &h, int p, double expected_inventory, const double *qbar, lsmc_trigger_prices &tp) {
  const StorageModel &st = *h.model;
  std::memset(&tp, 0, sizeof(tp));
  const double loss = st.loss_fraction(p) * expected_inventory;
  RateRange range = st.inject_withdraw_range(p, expected_inventory);
  double dset[LSMC_MAX_DECISIONS];
  const int D = bang_bang_decision_set(range, expected_inventory, loss, h.tables.space_lo[p + 1], h.tables.space_hi[p + 1], h.tolerance,
                                       h.extra, dset);
  const double df = h.settle_df[p];
  auto price_for = [&](double volume, double alt_cont, double alt_volume, double alt_cost, double alt_consumed) {
    const double cont = average_continuation(h, p, qbar, expected_inventory + volume - loss);
    const double cont_change = cont - alt_cont;
    const double excess = volume - alt_volume;
    const double cost_change = st.decision_cost_npv(p, volume) - alt_cost;
    const double consumed_change = st.decision_consumed(p, volume) - alt_consumed;
    return (cont_change - cost_change) / (df * (excess + consumed_change));
  };
  constexpr int NV = LSMC_NUM_TRIGGER_VOLUMES;
  const double max_inject = *std::max_element(dset, dset + D);
  if (max_inject > 0) {
    double alt = 0.0;
    bool have = false;
    for (int i = 0; i < D; ++i)
      if (dset[i] >= 0 && (!have || dset[i] < alt)) alt = dset[i], have = true;
    if (max_inject > alt) {
      const double alt_cont = average_continuation(h, p, qbar, expected_inventory + alt - loss);
      const double alt_cost = st.decision_cost_npv(p, alt), alt_cons = st.decision_consumed(p, alt);
      const double inc = (max_inject - alt) / NV;
      for (int i = 1; i < NV; ++i) tp.inject_volume[i - 1] = alt + i * inc;
      tp.inject_volume[NV - 1] = max_inject;
      for (int i = 0; i < NV; ++i) tp.inject_price[i] = price_for(tp.inject_volume[i], alt_cont, alt, alt_cost, alt_cons);
      tp.num_inject = NV;
      tp.has_inject_price = 1;
      tp.max_inject_trigger_price = tp.inject_price[NV - 1];
      tp.max_inject_volume = max_inject;
    }
  }
  const double max_withdraw = *std::min_element(dset, dset + D);
  if (max_withdraw < 0) {
    double alt = 0.0;
    bool have = false;
    for (int i = 0; i < D; ++i)
      if (dset[i] <= 0 && (!have || dset[i] > alt)) alt = dset[i], have = true;
    if (max_withdraw < alt) {
      const double alt_cont = average_continuation(h, p, qbar, expected_inventory + alt - loss);
      const double alt_cost = st.decision_cost_npv(p, alt), alt_cons = st.decision_consumed(p, alt);
      const double inc = (alt - max_withdraw) / NV;
      for (int i = 0; i < NV; ++i) {  // volumes ascend from the max withdrawal; the reference lists them reversed
        const double v = max_withdraw + (NV - 1 - i) * inc;
        tp.withdraw_volume[i] = v;
        tp.withdraw_price[i] = price_for(v, alt_cont, alt, alt_cost, alt_cons);
      }
      tp.num_withdraw = NV;
      tp.has_withdraw_price = 1;
      tp.max_withdraw_trigger_price = tp.withdraw_price[0];
      tp.max_withdraw_volume = max_withdraw;
    }
  }
}

void build_device_tables(lsmc_handle *h) {
  const int P = h->P, D = h->D, Gmax = h->Gmax;
  std::vector<DevDecision> dec;
  std::vector<int32_t> nd;
  std::vector<double> invc;
  h->table_offset.assign(P + 1, 0);
  size_t total = 0;
  for (int p = 0; p < P; ++p) {
    h->table_offset[p] = total;
    total += h->tables.periods[p].grid.size();
  }
  h->table_offset[P] = total;
  dec.resize(total * D);
  nd.resize(total);
  invc.resize(total);
  for (int p = 0; p < P; ++p) {
    const PeriodTables &pt = h->tables.periods[p];
    const size_t off = h->table_offset[p];
    for (size_t g = 0; g < pt.grid.size(); ++g) {
      nd[off + g] = pt.num_decisions[g];
      invc[off + g] = pt.inventory_cost[g];
      for (int d = 0; d < D; ++d) {
        const DecisionEntry &e = pt.decisions[g * D + d];
        dec[(off + g) * D + d] = DevDecision{e.volume, e.cost_npv, e.consumed, e.lower_w, e.upper_w, e.lower, e.upper};
      }
    }
  }
  h->decisions.upload(dec);
  h->num_decisions.upload(nd);
  h->inv_cost.upload(invc);

  // Row windows of the next period's value slab that each block of 8 grid points reads, for the row-ring sweep kernel.
  {
    constexpr int kGroup = 8;
    std::vector<SweepBlock> blocks_all;
    h->sweep_ring_rows.assign(P, 0);
    h->window_offset.assign(P, 0);
    for (int p = 0; p < P; ++p) {
      const PeriodTables &pt = h->tables.periods[p];
      const int G = (int)pt.grid.size(), groups = (G + kGroup - 1) / kGroup;
      h->window_offset[p] = blocks_all.size();
      std::vector<int32_t> lo(groups, INT32_MAX), hi(groups, -1);
      for (int g = 0; g < G; ++g)
        for (int d2 = 0; d2 < pt.num_decisions[g]; ++d2) {
          const DecisionEntry &e = pt.decisions[(size_t)g * D + d2];
          lo[g / kGroup] = std::min(lo[g / kGroup], e.lower);
          hi[g / kGroup] = std::max(hi[g / kGroup], e.upper);
        }
      // Row-ring sweep kernels (backward_sweep.cu, backward_sweep_mma.cu): while block j is evaluated the rows the next
      // `look_ahead` blocks add are copied in, so the ring must hold [lo(j), hi(j+look_ahead)] with hi a running maximum; usable
      // when lo never decreases and the ring fits beside the kernel's other shared-memory buffers with two CTAs per SM.
      {
        constexpr int kMaxSweepRing = 64;
        const int look_ahead = h->use_mma ? kMmaLookAhead : 1;  // blocks whose rows are requested ahead of the one being evaluated
        const char *env = std::getenv("LSMC_SWEEP_RING");  // test hook: 0 forces the direct-load sweep kernel
        bool sweep_ok = !(env && env[0] == '0');
        std::vector<int32_t> hic(hi);
        for (int j = 1; j < groups; ++j) hic[j] = std::max(hic[j], hic[j - 1]);
        int sweep_need = 0;
        for (int j = 0; j < groups; ++j) {
          if (j + 1 < groups && lo[j + 1] < lo[j]) sweep_ok = false;
          // (the FP64-mma kernel may still be reading block j - kMmaSlack when it requests the rows of block j + look_ahead)
          const int oldest = h->use_mma ? std::max(j - kMmaSlack, 0) : j;
          sweep_need = std::max(sweep_need, hic[std::min(j + look_ahead, groups - 1)] - lo[oldest] + 1);
        }
        if (sweep_need > kMaxSweepRing) sweep_ok = false;
        const int R = sweep_ok ? std::max(sweep_need, 2) : 0;
        h->sweep_ring_rows[p] = R;
        for (int j = 0; j < groups; ++j)
          blocks_all.push_back(SweepBlock{lo[j], hic[j], R ? lo[j] % R : 0, R ? hic[j] % R : 0});
      }
    }
    h->sweep_blocks.upload(blocks_all);
  }

  std::vector<double> grids((size_t)(P + 1) * Gmax, 0.0);
  for (int p = 0; p <= P; ++p) std::copy(h->tables.periods[p].grid.begin(), h->tables.periods[p].grid.end(), grids.begin() + (size_t)p * Gmax);
  h->grids.upload(grids);
  {  // which of mid = j and mid = j-1 the reference's bisection meets first, per next-period grid (SH.cs:280-314; bracket_dev)
    const size_t pitch = (size_t)forward_first_pitch(Gmax);
    std::vector<uint8_t> first((size_t)P * pitch, 1);
    for (int p = 0; p < P; ++p) {
      const std::vector<double> &g = h->tables.periods[p + 1].grid;
      for (int j = 1; j < (int)g.size(); ++j) {
        int lo = j, hi = j;
        bisect_inventory_space(g.data(), (int)g.size(), g[j], h->tolerance, &lo, &hi);
        first[(size_t)p * pitch + j] = lo == hi ? 1 : 0;
      }
    }
    h->bisect_first.upload(first);
  }

  const StorageModel &st = *h->model;
  std::vector<DevPeriodParams> pp(P);
  for (int p = 0; p < P; ++p) {
    DevPeriodParams &d = pp[p];
    d.min_inventory = st.min_inventory(p);
    d.max_inventory = st.max_inventory(p);
    d.inj_cost = st.inj_cost_[p];
    d.inj_df = st.inj_df_[p];
    d.wd_cost = st.wd_cost_[p];
    d.wd_df = st.wd_df_[p];
    d.cons_inj = st.cons_inj_[p];
    d.cons_wd = st.cons_wd_[p];
    d.loss = st.loss_[p];
    d.inv_cost = st.inv_cost_[p];
    d.inv_cost_df = st.inv_cost_df_[p];
    d.next_space_lo = h->tables.space_lo[p + 1];
    d.next_space_hi = h->tables.space_hi[p + 1];
    d.settle_df = h->settle_df[p];
    d.forward_price = h->forward_price[p];
    d.delta_df = h->discount_deltas ? h->settle_df[p] : 1.0;
    d.kind = st.kind_[p];
    d.rbegin = st.rbegin_[p];
    d.rcount = st.rcount_[p];
    d.poly_begin = st.poly_begin_[p] < 0 ? 0 : st.poly_begin_[p];
    d.grid_next_count = (int32_t)h->tables.periods[p + 1].grid.size();
    d.mean_mode = (p == 0 && h->current_active) ? 1 : 0;
    d.sim_row = d.mean_mode ? -1 : p - h->first_sim;
    d.pad = 0;
  }
  h->period_params.upload(pp);
  h->ratchet_inv.upload(st.r_inv_);
  h->ratchet_min.upload(st.r_min_);
  h->ratchet_max.upload(st.r_max_);
  {  // spline slopes per ratchet segment (the last knot of a period has none)
    std::vector<double> smin(st.r_inv_.size(), 0.0), smax(st.r_inv_.size(), 0.0);
    for (int p = 0; p < P; ++p)
      for (int k = 0; k + 1 < st.rcount_[p]; ++k) {
        const size_t i = (size_t)st.rbegin_[p] + k;
        const double dx = st.r_inv_[i + 1] - st.r_inv_[i];
        smin[i] = (st.r_min_[i + 1] - st.r_min_[i]) / dx;
        smax[i] = (st.r_max_[i + 1] - st.r_max_[i]) / dx;
      }
    h->ratchet_min_slope.upload(smin);
    h->ratchet_max_slope.upload(smax);
  }
  h->poly_min.upload(st.poly_min_);
  h->poly_max.upload(st.poly_max_);
  h->d_powers.upload(h->powers);
}

void validate_problem(const lsmc_problem *pb) {
  if (!pb) throw Error(LSMC_ERR_ARGUMENT, "problem is null");
  if (pb->inventory < 0) throw Error(LSMC_ERR_ARGUMENT, "Inventory cannot be negative.");  // LSV.cs:64-65
  if (pb->storage.num_periods < 1) throw Error(LSMC_ERR_ARGUMENT, "Storage has expired");
  if (pb->num_sims <= 0) throw Error(LSMC_ERR_ARGUMENT, "Number of simulations must be positive.");
  if (pb->num_basis < 1 || pb->num_basis > LSMC_MAX_BASIS) throw Error(LSMC_ERR_ARGUMENT, "Number of basis functions must be in [1, LSMC_MAX_BASIS].");
  if (pb->num_factors < 0 || pb->num_factors > LSMC_MAX_FACTORS) throw Error(LSMC_ERR_ARGUMENT, "Number of factors must be in [0, LSMC_MAX_FACTORS].");
  if (pb->extra_decisions < 0) throw Error(LSMC_ERR_INVALID_OPERATION, "ExtraDecisions must be non-negative.");  // LVP.cs:137-138
  if (2 * pb->extra_decisions + 3 > LSMC_MAX_DECISIONS) throw Error(LSMC_ERR_ARGUMENT, "ExtraDecisions too large for LSMC_MAX_DECISIONS.");
  if (!(pb->grid_spacing > 0.0)) throw Error(LSMC_ERR_ARGUMENT, "Grid spacing must be positive.");
  if (!pb->forward_price || !pb->settle_df || !pb->basis_powers) throw Error(LSMC_ERR_INVALID_OPERATION, "ForwardCurve, SettleDateRule and BasisFunctions must be set.");
  for (int i = 0; i < pb->num_basis * (1 + pb->num_factors); ++i)
    if (pb->basis_powers[i] < 0 || pb->basis_powers[i] > 64) throw Error(LSMC_ERR_ARGUMENT, "Spot power and Markov factor powers must be non-negative.");
}

void run_valuation(lsmc_handle *h, lsmc_progress_fn progress, void *user, const volatile int *cancel) {
  CUDA_OK(cudaSetDevice(h->device));
  if (!h->sims[0].ready) throw Error(LSMC_ERR_INVALID_OPERATION, "RegressionSpotSimsGenerator has not been set.");
  if (!h->sims[1].ready) throw Error(LSMC_ERR_INVALID_OPERATION, "ValuationSpotSimsGenerator has not been set.");
  const int P = h->P, K = h->K, F = h->F, D = h->D, Gmax = h->Gmax;
  const int64_t N = h->N;
  cudaStream_t st = h->stream;
  h->sims[0].wait_upload();  // the regression simulations are read from the first kernel on
  if (h->sims[0].upload_stream) CUDA_OK(cudaStreamSynchronize(h->sims[0].upload_stream));
  h->has_results = false;
  h->scalars = lsmc_results{};
  h->scalars.ms_regression_sim = h->sims[0].ms_simulation;
  h->scalars.ms_valuation_sim = h->sims[1].ms_simulation;
  int64_t &launches = h->scalars.kernel_launches;
  auto wall0 = std::chrono::steady_clock::now();
  // status block: [0] error flag (an lsmc_status), [2..3] 64-bit count of inventories moved onto a limit
  CUDA_OK(cudaMemsetAsync(h->error_flag.ptr, 0, 4 * sizeof(int32_t), st));
  auto cancelled = [&] { return cancel && *cancel != 0; };
  // Sharded runs honour a cancellation after the device work, when the ranks exchange their status: a rank that stopped
  // enqueueing on its own would leave the others waiting for it in the next per-step exchange.
  auto check_cancel = [&] {
    if (cancelled() && h->comm.world() == 1) {
      cudaStreamSynchronize(st);
      cudaStreamSynchronize(h->aux_stream);
      throw Error(LSMC_ERR_CANCELLED, "The operation was canceled.");
    }
  };
  check_cancel();

  const bool debug_sync = std::getenv("LSMC_DEBUG_SYNC") != nullptr;
  auto dbg = [&](const char *what, int p) {
    if (!debug_sync) return;
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) throw CudaFailure(std::string(what) + " (period " + std::to_string(p) + ") failed: " + cudaGetErrorString(e));
  };
  Timer t_pinv(st), t_back(st), t_fwd(st), t_dev(st);
  struct EventList {
    std::vector<cudaEvent_t> ev;
    explicit EventList(size_t n) : ev(n, nullptr) {
      for (auto &e : ev) CUDA_OK(cudaEventCreate(&e));
    }
    ~EventList() {
      for (auto &e : ev)
        if (e) cudaEventDestroy(e);
    }
    cudaEvent_t &operator[](size_t i) { return ev[i]; }
  };
  std::vector<cudaEvent_t> &step_events = h->step_events;
  while (step_events.size() < 2 * (size_t)P) {
    cudaEvent_t e = nullptr;
    CUDA_OK(cudaEventCreate(&e));
    step_events.push_back(e);
  }
  t_dev.start();
  // ---- regression pre-pass: Gram matrices, Cholesky factors and R^-1 of every regression period (LSV.cs:191-196)
  const int first_reg = h->current_active ? 1 : 0;
  const int num_slots = P - first_reg;
  t_pinv.start();
  if (num_slots > 0) {
    const int NP = h->KP * (h->KP + 1) / 2;
    int64_t chunk_paths = ((N + 63) / 64 + kThreads - 1) / kThreads * kThreads;
    chunk_paths = std::max<int64_t>(chunk_paths, 4 * kThreads);
    const int num_chunks = (int)((N + chunk_paths - 1) / chunk_paths);
    const int world = h->comm.world();
    std::vector<BasisSpec> specs(num_slots);
    for (int s = 0; s < num_slots; ++s) specs[s] = basis_spec(*h, 0, first_reg + s);
    h->gram_specs.upload(specs);
    // every rank's partial Gram matrices, gathered as [rank][slot][chunk][NP][2]; the Cholesky kernel sums them in
    // (rank, chunk) order so all ranks obtain bit-identical factors
    const size_t local_count = (size_t)num_slots * num_chunks * NP * 2;
    h->gram_partials.allocate(local_count * world);
    GramArgs ga{};
    ga.N = N;
    ga.chunk_paths = chunk_paths;
    ga.K = K;
    ga.F = F;
    ga.specs = h->gram_specs.ptr;
    ga.powers = h->d_powers.ptr;
    ga.partials = h->gram_partials.ptr + (size_t)h->comm.rank() * local_count;
    launch_gram(ga, num_chunks, num_slots, st);
    launches++;
    if (world > 1) h->comm.all_gather(ga.partials, h->gram_partials.ptr, local_count, st);
    CholeskyArgs ca{};
    ca.K = K;
    ca.num_ranks = world;
    ca.num_slots = num_slots;
    ca.num_chunks = num_chunks;
    ca.partials = h->gram_partials.ptr;
    ca.rinv = h->rinv.ptr + (size_t)first_reg * K * K * 2;
    ca.error_flag = h->error_flag.ptr;
    dbg("gram_kernel", -1);
    launch_cholesky(ca, num_slots, st);
    launches++;
    dbg("cholesky_kernel", -1);
  }
  t_pinv.stop();

  // ---- backward induction (LSV.cs:111-131, 169-346)
  t_back.start();
  const int64_t num_tiles = (N + kTilePaths - 1) / kTilePaths;
  int sm_count = 148;
  CUDA_OK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, h->device));
  // The dominant kernel is timed with CUDA events around its launch.  An event record between two dependent kernels costs the
  // GPU front end a little each time, which matters when a step is ~100 us (sharded runs, small configurations): long runs time
  // every kStepSample-th step (the grid width varies slowly with the period, so the sample is representative), LSMC_STEP_EVENTS=all
  // times every step, LSMC_NO_STEP_EVENTS none.
  const bool time_steps = std::getenv("LSMC_NO_STEP_EVENTS") == nullptr;
  const int kStepSample = [&] {
    const char *e = std::getenv("LSMC_STEP_EVENTS");
    return (e && e[0] == 'a') || P <= 64 ? 1 : 8;
  }();
  auto timed_step = [&](int p) { return time_steps && p % kStepSample == 0; };
  const bool trace = std::getenv("LSMC_TRACE") != nullptr;
  EventList tr(trace ? 5 * (size_t)(P + 1) : 0);
  auto mark = [&](int p, int i) { if (trace) CUDA_OK(cudaEventRecord(tr[5 * (size_t)p + i], st)); };
  // a launch that could not be enqueued (bad configuration, sticky error) is reported with its kernel and period
  auto launched = [&](const char *what, int p) {
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) throw CudaFailure(std::string(what) + " (period " + std::to_string(p) + ") could not be launched: " + cudaGetErrorString(e));
  };
  const bool use_sweep = [] { const char *e = std::getenv("LSMC_SWEEP"); return !(e && e[0] == '0'); }();  // test hook: 0 = general kernel
  auto q_buffer = [&](int period) { return h->qbuf[period % 3].ptr; };
  auto has_basis = [&](int period) { return period >= first_reg && period < P; };
  auto launch_basis_for = [&](int period) {
    // orthonormal basis rows of the regression simulations of `period` (LSV.cs:191 PopulateDesignMatrix + :193-196 QR)
    BasisArgs b{};
    b.N = N;
    b.pitch = h->pitch;
    b.K = K;
    b.F = F;
    b.spec = basis_spec(*h, 0, period);
    b.powers = h->d_powers.ptr;
    b.q_out = q_buffer(period);
    launch_basis(b, h->aux_stream);
    launched("basis_kernel", period);
    CUDA_OK(cudaEventRecord(h->ev_basis[period % 3], h->aux_stream));
    launches++;
    if (debug_sync) {
      cudaError_t e = cudaStreamSynchronize(h->aux_stream);
      if (e == cudaSuccess) e = cudaGetLastError();
      if (e != cudaSuccess) throw CudaFailure(std::string("basis_kernel (period ") + std::to_string(period) + ") failed: " + cudaGetErrorString(e));
    }
  };
  // The basis rows are independent of the value recursion, so they are produced on a second stream, three buffers deep
  // (equal stream priorities: measured, a low-priority basis stream ends up sharing the GPU with the sweep kernel and costs
  // more than it hides; as it is, the basis kernel fills the SMs the small reduction and fold kernels leave idle):
  // step p reads q(p) and q(p-1); once it has finished, the buffer of q(p) is free for q(p-3), which is then computed while
  // step p-1 runs and is first needed by step p-2.
  CUDA_OK(cudaEventRecord(h->ev_prepass, st));
  CUDA_OK(cudaStreamWaitEvent(h->aux_stream, h->ev_prepass, 0));
  if (has_basis(P - 1)) launch_basis_for(P - 1);
  if (has_basis(P - 2)) launch_basis_for(P - 2);
  auto uses_sweep = [&](int p) { return D == 3 && use_sweep; };
  // per-grid-point records of step p: payoffs, weights, row offsets and the folded regression coefficients (needs gamma[p])
  auto fold_args_for = [&](int p) {
    const PeriodTables &pt = h->tables.periods[p];
    FoldArgs f{};
    f.G = (int)pt.grid.size();
    f.Gn = (int)h->tables.periods[p + 1].grid.size();
    f.K = K;
    f.pitch = h->pitch;
    f.decisions = h->decisions.ptr + h->table_offset[p] * D;
    f.num_decisions = h->num_decisions.ptr + h->table_offset[p];
    f.inventory_cost = h->inv_cost.ptr + h->table_offset[p];
    f.gamma = h->gamma.ptr + (size_t)p * Gmax * K;
    f.settle_df = h->settle_df[p];
    f.fused_payoff = (h->flags & LSMC_REFERENCE_OPERATION_ORDER) ? 0 : 1;
    f.ring_rows = h->sweep_ring_rows[p];
    f.format = (h->use_mma && f.ring_rows > 0) ? 1 : 0;
    f.row_pitch_bytes = kMmaRowPitchBytes;
    f.out = h->point_tables.ptr;
    return f;
  };
  struct StepSums {  // where a step left its regression partial sums, and how many chunks per grid point
    const double *partials;
    int chunks;
  };
  auto launch_step = [&](int p, bool terminal, const double *v_next, double *v_this, bool fold_done) {
    const PeriodTables &pt = h->tables.periods[p];
    BackwardStepArgs a{};
    a.N = N;
    a.pitch = h->pitch;
    a.G = (int)pt.grid.size();
    a.Gn = terminal ? 0 : (int)h->tables.periods[p + 1].grid.size();
    a.D = terminal ? 3 : D;
    a.K = K;
    a.v_next = v_next;
    a.v_this = v_this;
    a.terminal_mode = terminal ? 1 : 0;
    a.fused_payoff = (h->flags & LSMC_REFERENCE_OPERATION_ORDER) ? 0 : 1;
    if (terminal) {
      a.grid_this = h->grids.ptr + (size_t)p * Gmax;
      a.spot = h->sims[0].spot + (size_t)(p - h->first_sim) * N;
      std::memcpy(a.terminal, h->model->terminal_, sizeof(a.terminal));
    } else {
      a.gamma = h->gamma.ptr + (size_t)p * Gmax * K;
      a.decisions = h->decisions.ptr + h->table_offset[p] * D;
      a.num_decisions = h->num_decisions.ptr + h->table_offset[p];
      a.inventory_cost = h->inv_cost.ptr + h->table_offset[p];
      a.settle_df = h->settle_df[p];
      if (has_basis(p)) {
        a.q_this = q_buffer(p);
        a.spot = h->sims[0].spot + (size_t)(p - h->first_sim) * N;
      } else {  // current period: deterministic price, sample-mean continuation value (LSV.cs:174-185,232-236)
        a.q_this = nullptr;
        a.spot = nullptr;
        a.const_price = h->forward_price[p];
      }
    }
    a.accumulate = p > 0 ? 1 : 0;
    a.q_prev = has_basis(p - 1) ? q_buffer(p - 1) : nullptr;
    a.partials = h->partials.ptr;
    // split the inventory grid over several CTAs only when there are too few path tiles to fill the GPU
    const bool sweep = !terminal && uses_sweep(p);
    const int sub_size = sweep ? 8 : kMaxGridTile;
    const int subtiles = (a.G + sub_size - 1) / sub_size;
    const int64_t tiles = sweep ? (N + kSweepTilePaths - 1) / kSweepTilePaths : num_tiles;
    const int64_t slots = (int64_t)sm_count * (sweep ? 3 : 4);
    int gsplit = (int)std::min<int64_t>(subtiles, std::max<int64_t>(1, (4 * slots + tiles - 1) / tiles));
    if (const char *forced = std::getenv("LSMC_GSPLIT")) gsplit = std::max(1, std::min(subtiles, std::atoi(forced)));  // test hook
    a.subtiles_per_cta = (subtiles + gsplit - 1) / gsplit;
    gsplit = (subtiles + a.subtiles_per_cta - 1) / a.subtiles_per_cta;
    if (sweep) {
      if (!fold_done) {
        const FoldArgs f = fold_args_for(p);
        mark(p, 1);
        launch_fold_tables(f, st);
        mark(p, 2);
        launches++;
      }
      a.point_tables = h->point_tables.ptr;
      a.ring_rows = h->sweep_ring_rows[p];
      a.sweep_blocks = h->sweep_blocks.ptr + h->window_offset[p];
    }
    mark(p, 3);
    if (!terminal && timed_step(p)) CUDA_OK(cudaEventRecord(step_events[2 * p], st));
    StepSums out{h->partials.ptr, (int)tiles};
    if (sweep && h->use_mma && a.ring_rows > 0) {
      // Whole tiles for as many full waves of resident CTA slots as there are; the tiles of the last, partial wave are split
      // into the number of pieces that minimises (waves of pieces) x (blocks per piece + one block-time of start-up).
      const int64_t slots = (int64_t)sm_count * backward_sweep_mma_ctas_per_sm();
      const int blocks = (a.G + 7) / 8;
      a.num_tiles = tiles;
      a.whole_tiles = tiles / slots * slots;
      a.tail_split = 1;
      const int64_t rest = tiles - a.whole_tiles;
      if (rest > 0) {
        int64_t best = INT64_MAX;
        for (int g = 1; g <= std::max(1, std::min(12, blocks / 2)); ++g) {
          const int64_t cost = ((rest * g + slots - 1) / slots) * ((blocks + g - 1) / g + 1);
          if (cost < best) best = cost, a.tail_split = g;
        }
      }
      if (const char *forced = std::getenv("LSMC_MMA_TAIL_SPLIT")) {  // test hook: "<whole tiles>,<pieces per remaining tile>"
        a.whole_tiles = std::max<int64_t>(0, std::min<int64_t>(tiles, std::atoll(forced)));
        const char *comma = std::strchr(forced, ',');
        a.tail_split = std::max(1, std::min(blocks, comma ? std::atoi(comma + 1) : 1));
      }
      launch_backward_sweep_mma(a, st);
    } else if (sweep) {
      launch_backward_sweep(a, gsplit, tiles, st);
    } else {
      launch_backward_step(a, gsplit, num_tiles, st);
    }
    launched(sweep ? "backward sweep kernel" : "backward_step_kernel", p);
    if (!terminal && timed_step(p)) CUDA_OK(cudaEventRecord(step_events[2 * p + 1], st));
    mark(p, 4);
    launches++;
    dbg(sweep ? "backward_sweep_kernel" : "backward_step_kernel", p);
    if (has_basis(p - 3)) {
      CUDA_OK(cudaEventRecord(h->ev_step[p % 3], st));
      CUDA_OK(cudaStreamWaitEvent(h->aux_stream, h->ev_step[p % 3], 0));
      launch_basis_for(p - 3);
    }
    return out;
  };
  auto reduce_gamma = [&](int p, int G_next, StepSums sums) -> bool {  // true: the point tables of step p were folded as well
    // regression coefficients of V_{p+1} on period p's basis (LSV.cs:199-212) from the partial sums the kernel that
    // produced V_{p+1} left behind
    const bool mean_mode = (p == 0 && h->current_active);
    ReduceArgs r{};
    r.partials = sums.partials;
    r.num_chunks = sums.chunks;
    r.G = G_next;
    r.K = K;
    r.acc_k = h->acc_k;
    r.mean_mode = mean_mode ? 1 : 0;
    r.total_paths = h->comm.total_sims(N);
    double *dst = h->gamma.ptr + (size_t)p * Gmax * K;
    const int world = h->comm.world();
    if (world == 1) {
      r.out = dst;
      if (uses_sweep(p)) {  // reduction and table fold in one launch
        launch_reduce_fold(r, fold_args_for(p), h->fold_counter.ptr, st);
        launches++;
        return true;
      }
      launch_reduce_partials(r, st);
      launches++;
    } else if (h->comm.peer_exchange_ready()) {
      // local sums -> every rank's inbox over peer memory -> sum in rank order -> table fold, in one launch
      r.out = dst;
      FoldArgs f{};
      if (uses_sweep(p)) f = fold_args_for(p);
      launch_exchange_fold(r, f, h->comm.next_exchange(h->error_flag.ptr), h->fold_counter.ptr, st);
      launches++;
      return uses_sweep(p);
    } else {
      // local sums -> all-gather -> sum in rank order (identical on every rank)
      const int count = G_next * K;
      if (h->gather_buf.count < (size_t)count * world) h->gather_buf.allocate((size_t)Gmax * K * world);
      r.mean_mode = 0;
      r.out = h->gather_buf.ptr + (size_t)h->comm.rank() * count;
      launch_reduce_partials(r, st);
      h->comm.all_gather(r.out, h->gather_buf.ptr, count, st);
      launch_combine_ranks(h->gather_buf.ptr, world, count, K, mean_mode ? 1 : 0, (double)r.total_paths, dst, st);
      launches += 2;
    }
    return false;
  };

  double *v_next = h->v_a.ptr, *v_this = h->v_b.ptr;
  const auto enqueue0 = std::chrono::steady_clock::now();
  // the basis rows a step needs (period p - 1, for the regression sums) are waited for BEFORE the step's reduction kernel, so that
  // nothing stands between that kernel and the sweep kernel that may overlap its tail (programmatic dependent launch)
  auto wait_basis_for_step = [&](int p) {
    if (has_basis(p - 1)) CUDA_OK(cudaStreamWaitEvent(st, h->ev_basis[(p - 1) % 3], 0));
  };
  wait_basis_for_step(P);
  StepSums chunks = launch_step(P, true, nullptr, v_next, false);
  const double back_step = 0.66 / P;  // BackwardPcntTime (LSV.cs:48,164)
  double prog = 0.0;
  // Steps are enqueued in windows; a window's progress notifications (one per step, as the reference gives them) and the
  // cancellation poll happen when the GPU has FINISHED that window, two windows behind the enqueueing, so that the GPU always
  // has at least one whole window queued while the host waits.
  constexpr int kWindow = 64, kWindowsInFlight = 2;
  int windows_enqueued = 0, windows_reported = 0;
  auto report_window = [&](int w) {  // blocks until window w is done on the GPU
    CUDA_OK(cudaEventSynchronize(h->ev_window[w % 3]));
    const int first = P - 1 - w * kWindow, last = std::max(first - kWindow + 1, 0);
    for (int q = first; q >= last; --q) {
      prog += back_step;
      if (progress) progress(prog, user);
    }
    check_cancel();
  };
  for (int p = P - 1; p >= 0; --p) {
    mark(p, 0);
    wait_basis_for_step(p);
    const bool folded = reduce_gamma(p, (int)h->tables.periods[p + 1].grid.size(), chunks);
    launched("per-step reduction", p);
    chunks = launch_step(p, false, v_next, v_this, folded);
    std::swap(v_next, v_this);
    if ((P - 1 - p) % kWindow == kWindow - 1 || p == 0) {
      CUDA_OK(cudaEventRecord(h->ev_window[windows_enqueued % 3], st));
      ++windows_enqueued;
      while (windows_enqueued - windows_reported > kWindowsInFlight) report_window(windows_reported++);
    }
  }
  const double *v_first = v_next;  // storage value by sim at period 0
  if (trace) {
    cudaStreamSynchronize(st);
    for (int p = P / 2 + 2; p >= P / 2; --p) {
      float m[5];
      for (int i = 0; i < 5; ++i)
        if (cudaEventElapsedTime(&m[i], tr[5 * (size_t)(P / 2 + 2)], tr[5 * (size_t)p + i]) != cudaSuccess) {
          m[i] = 0.f;  // a mark this code path never recorded (reduction and fold share one launch)
          cudaGetLastError();
        }
      std::fprintf(stderr, "[lsmc] step %d: start %.1f us | reduce done %.1f | fold done %.1f | sweep may start %.1f | sweep done %.1f\n", p,
                   1e3 * m[0], 1e3 * m[1], 1e3 * m[2], 1e3 * m[3], 1e3 * m[4]);
    }
  }
  if (trace)
    std::fprintf(stderr, "[lsmc] backward induction: %d steps enqueued in %.3f ms of host time\n", P,
                 std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - enqueue0).count());
  t_back.stop();

  // ---- forward simulation (LSV.cs:390-521)
  t_fwd.start();
  const int sums_width = kForwardSums + K;
  const int64_t fwd_tiles = (N + kThreads - 1) / kThreads;
  constexpr int kForwardMaxTiles = 8;  // per-path state of this many tiles fits beside the period tables in shared memory
  // The kernel walks all periods, so its CTAs should all be resident at once (4 per SM) or form whole waves of that many: with
  // 1.65 waves the second wave costs as much as the first.  Tiles are dealt round-robin over the CTAs (tile t -> CTA t mod
  // ctas), which spreads the CTAs that carry one tile more than the others evenly over the SMs.
  const int64_t fwd_slots = (int64_t)sm_count * 4;
  const int64_t fwd_waves = (fwd_tiles + fwd_slots * kForwardMaxTiles - 1) / (fwd_slots * kForwardMaxTiles);
  const int fwd_ctas = (int)std::min<int64_t>(fwd_tiles, fwd_slots * std::max<int64_t>(1, fwd_waves));
  if (h->step_sums.count < (size_t)P * fwd_ctas * sums_width) h->step_sums.allocate((size_t)P * fwd_ctas * sums_width);
  if (h->step_totals.count < (size_t)P * sums_width) h->step_totals.allocate((size_t)P * sums_width);
  ForwardArgs fa{};
  fa.N = N;
  fa.P = P;
  fa.K = K;
  fa.F = F;
  fa.Gmax = Gmax;
  fa.sim_pitch = N;
  fa.periods = h->period_params.ptr;
  fa.ratchet_inventory = h->ratchet_inv.ptr;
  fa.ratchet_min = h->ratchet_min.ptr;
  fa.ratchet_max = h->ratchet_max.ptr;
  fa.ratchet_min_slope = h->ratchet_min_slope.ptr;
  fa.ratchet_max_slope = h->ratchet_max_slope.ptr;
  fa.poly_min = h->poly_min.ptr;
  fa.poly_max = h->poly_max.ptr;
  fa.grids = h->grids.ptr;
  fa.bisect_first = h->bisect_first.ptr;
  fa.first_pitch = forward_first_pitch(Gmax);
  fa.gamma = h->gamma.ptr;
  fa.rinv = h->rinv.ptr;
  fa.powers = h->d_powers.ptr;
  fa.spot = h->sims[1].spot;
  for (int i = 0; i < F; ++i) fa.factors[i] = h->sims[1].factors[i];
  fa.start_inventory = h->inventory;
  fa.tolerance = h->tolerance;
  fa.inv_grid_spacing = 1.0 / h->grid_spacing;
  fa.extra_decisions = h->extra;
  fa.pv_by_sim = h->pv.ptr;
  fa.final_inventory = h->final_inv.ptr;
  fa.step_sums = h->step_sums.ptr;
  fa.sums_width = sums_width;
  fa.tiles_per_cta = (int)((fwd_tiles + fwd_ctas - 1) / fwd_ctas);
  // Two table buffers (the next period's tables fetched under this period's arithmetic) when a CTA carries one or two tiles and
  // has nothing else to hide the fetch behind: cfg 1 forward pass -6 %, cfg 2 -3 %.  With many tiles per CTA the other CTAs of
  // the SM cover it already and the second buffer only costs (1M paths: +3 %), so one buffer is used there.
  fa.table_buffers = fa.tiles_per_cta <= 2 && forward_smem_bytes(K, F, Gmax, fa.first_pitch, 2, sums_width, fa.tiles_per_cta) <= 54 * 1024 ? 2 : 1;
  if (const char *e = std::getenv("LSMC_FWD_TABLE_BUFFERS")) fa.table_buffers = std::atoi(e) >= 2 ? 2 : 1;  // test hook
  fa.error_flag = h->error_flag.ptr;
  fa.snapped = reinterpret_cast<unsigned long long *>(h->error_flag.ptr + 2);
  // the valuation simulations may still be on their way (lsmc_set_sims_async): the forward pass is their first reader
  h->sims[1].wait_upload();
  if (h->sims[1].upload_stream) CUDA_OK(cudaStreamSynchronize(h->sims[1].upload_stream));
  double **panel_arg[7] = {&fa.p_inventory, &fa.p_volume, &fa.p_consumed, &fa.p_loss, &fa.p_net_volume, &fa.p_pv, &fa.p_decision};
  if (!h->stream_panels) {
    for (int i = 0; i < 7; ++i) *panel_arg[i] = h->panels[i].ptr;
    fa.p_begin = 0;
    fa.p_end = P;
    fa.panel_row0 = 0;
    launch_forward(fa, fwd_ctas, st);
    launched("forward_kernel", -1);
    launches++;
    // the forward pass is queued behind the last windows: report those while it runs
    while (windows_reported < windows_enqueued) report_window(windows_reported++);
  } else {
    // Streamed panels: blocks of `stream_block` periods alternate between the two halves of every panel ring;
    // while block b is simulated the host drains block b-1 through the pinned staging buffers into the caller's arrays.
    for (int i = 0; i < 7; ++i)
      if (h->panels[i].ptr && !h->sinks[i])
        throw Error(LSMC_ERR_INVALID_OPERATION, "LSMC_STREAM_PANELS: lsmc_set_panel_sink has not been called for every requested panel.");
    const int B = h->stream_block, num_blocks = (P + B - 1) / B;
    auto drain = [&](int b) {  // copies block b's rows of every panel to its sink; blocks the host until done
      const int first = b * B, rows = std::min(B, P - first);
      CUDA_OK(cudaStreamWaitEvent(h->copy_stream, h->ev_block[b & 1], 0));
      for (int i = 0; i < 7; ++i)
        if (h->panels[i].ptr)
          h->sink_stager.to_host(h->sinks[i] + (size_t)first * N, h->panels[i].ptr + (size_t)(b & 1) * B * N, (size_t)rows * N * sizeof(double),
                                 h->copy_stream);
      CUDA_OK(cudaEventRecord(h->ev_copied[b & 1], h->copy_stream));
    };
    for (int b = 0; b < num_blocks; ++b) {
      if (b >= 2) CUDA_OK(cudaStreamWaitEvent(st, h->ev_copied[b & 1], 0));  // the half this block writes has been drained
      for (int i = 0; i < 7; ++i) *panel_arg[i] = h->panels[i].ptr ? h->panels[i].ptr + (size_t)(b & 1) * B * N : nullptr;
      fa.p_begin = b * B;
      fa.p_end = std::min(P, (b + 1) * B);
      fa.panel_row0 = fa.p_begin;
      launch_forward(fa, fwd_ctas, st);
      launched("forward_kernel", fa.p_begin);
      launches++;
      CUDA_OK(cudaEventRecord(h->ev_block[b & 1], st));
      if (b == 0)
        while (windows_reported < windows_enqueued) report_window(windows_reported++);
      if (b >= 1) drain(b - 1);
    }
    drain(num_blocks - 1);
  }
  dbg("forward_kernel", -1);
  launch_reduce_step_sums(h->step_sums.ptr, P, fwd_ctas, sums_width, h->step_totals.ptr, st);
  launches++;
  t_fwd.stop();
  t_dev.stop();

  std::vector<double> sums((size_t)P * sums_width);
  CUDA_OK(cudaMemcpyAsync(sums.data(), h->step_totals.ptr, sums.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  h->h_gamma.resize((size_t)P * Gmax * K);
  CUDA_OK(cudaMemcpyAsync(h->h_gamma.data(), h->gamma.ptr, h->h_gamma.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  h->h_rinv.resize((size_t)P * K * K * 2);
  CUDA_OK(cudaMemcpyAsync(h->h_rinv.data(), h->rinv.ptr, h->h_rinv.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  int32_t status_block[4] = {0, 0, 0, 0};
  CUDA_OK(cudaMemcpyAsync(status_block, h->error_flag.ptr, sizeof(status_block), cudaMemcpyDeviceToHost, st));
  CUDA_OK(cudaStreamSynchronize(st));
  CUDA_OK(cudaGetLastError());
  int32_t flag = status_block[0];
  int64_t snapped = 0;
  std::memcpy(&snapped, &status_block[2], sizeof(snapped));
  if (h->comm.world() > 1) {
    // every rank must leave lsmc_run with the same status, or the ranks that carry on would wait for the one that threw in
    // the next collective: exchange (status, snapped count, cancel seen) before acting on any of them
    const double mine[3] = {(double)flag, (double)snapped, cancelled() ? 1.0 : 0.0};
    std::vector<double> all;
    h->comm.all_gather_host(mine, 3, all, st);
    flag = 0;
    snapped = 0;
    bool any_cancel = false;
    for (int r = 0; r < h->comm.world(); ++r) {
      if (flag == 0) flag = (int32_t)all[3 * r];  // first failing rank, in rank order
      snapped += (int64_t)all[3 * r + 1];
      any_cancel = any_cancel || all[3 * r + 2] != 0.0;
    }
    if (any_cancel) throw Error(LSMC_ERR_CANCELLED, "The operation was canceled.");
  }
  if (flag == LSMC_ERR_CUDA)
    throw Error(LSMC_ERR_CUDA, "A rank of the sharded valuation did not arrive at a per-step exchange within the time limit (it failed, was cancelled or is not running the same valuation).");
  if (flag == LSMC_ERR_NUMERICAL)
    throw Error(LSMC_ERR_NUMERICAL, "Regression design matrix is rank deficient on the simulated paths (the reference's QR would return non-finite coefficients).");
  if (flag != 0)
    throw Error((lsmc_status)flag, "A simulated path left the feasible inventory region during the forward simulation (inventory outside its limits or outside the inventory grid).");
  h->comm.all_reduce_host(sums.data(), sums.size(), st);
  const double total_n = (double)h->comm.total_sims(N);

  // per-period results (LSV.cs:510-518) and trigger prices (LSV.cs:523-592)
  h->deltas.assign(P + 1, 0.0);
  h->profile.assign((size_t)(P + 1) * 5, 0.0);
  h->trigger.assign(P, lsmc_trigger_prices{});
  const double fwd_step = (1.0 - 0.66) / (P + 1);  // LSV.cs:387
  for (int p = 0; p < P; ++p) {
    const double *s = &sums[(size_t)p * sums_width];
    double *pr = &h->profile[(size_t)p * 5];
    const double expected_inventory = s[5] / total_n;
    pr[0] = expected_inventory;
    pr[1] = s[1] / total_n;
    pr[2] = s[2] / total_n;
    pr[3] = s[3] / total_n;
    pr[4] = s[4] / total_n;
    h->deltas[p] = (s[0] / h->forward_price[p] / total_n) * (h->discount_deltas ? h->settle_df[p] : 1.0);
    double qbar[LSMC_MAX_BASIS];
    for (int k = 0; k < K; ++k) qbar[k] = s[kForwardSums + k] / total_n;
    trigger_prices_for_period(*h, p, expected_inventory, qbar, h->trigger[p]);
    prog += fwd_step;
    if (progress) progress(prog, user);
    check_cancel();
  }

  if (h->stream_panels) {  // row P of the streamed [P+1][N] panels: zero, except the inventory the paths end with (LSV.cs:627-629)
    for (int i = 0; i < 6; ++i)
      if (h->sinks[i] && h->panels[i].ptr) std::fill(h->sinks[i] + (size_t)P * N, h->sinks[i] + (size_t)(P + 1) * N, 0.0);
    if (h->sinks[0] && h->panels[0].ptr)
      CUDA_OK(cudaMemcpy(h->sinks[0] + (size_t)P * N, h->final_inv.ptr, N * sizeof(double), cudaMemcpyDeviceToHost));
  }

  // ---- terminal PV, NPV and standard error (LSV.cs:595-618)
  double end_period_pv = 0.0;
  const bool need_pv_host = !h->model->must_be_empty_at_end();
  h->pv_by_sim.clear();
  double npv, variance;
  if (need_pv_host) {
    std::vector<double> pv(N), inv(N), s_end(N);
    CUDA_OK(cudaMemcpy(pv.data(), h->pv.ptr, N * sizeof(double), cudaMemcpyDeviceToHost));
    CUDA_OK(cudaMemcpy(inv.data(), h->final_inv.ptr, N * sizeof(double), cudaMemcpyDeviceToHost));
    CUDA_OK(cudaMemcpy(s_end.data(), h->sims[0].spot + (size_t)(P - h->first_sim) * N, N * sizeof(double), cudaMemcpyDeviceToHost));  // regression sims
    std::vector<double> last_row(h->panels[5].ptr ? N : 0);
    // Sharded: the reference's running sum runs over the paths in global order, so a rank starts from the total of
    // the ranks before it (exclusive prefix of the per-rank totals, one scalar gathered per rank).
    double local_total = 0.0;
    for (int64_t n = 0; n < N; ++n) {
      s_end[n] = h->model->terminal_npv(s_end[n], inv[n]);  // terminal value of path n, in place
      local_total += s_end[n];
    }
    std::vector<double> totals;
    h->comm.all_gather_host(&local_total, 1, totals, st);
    double running = 0.0, global_total = 0.0;
    for (int r = 0; r < h->comm.world(); ++r) {
      if (r < h->comm.rank()) running += totals[r];
      global_total += totals[r];
    }
    for (int64_t n = 0; n < N; ++n) {
      const double t = s_end[n];
      running += t;
      const double add = h->compat_terminal ? running : t;  // cumulative-sum quirk of LSV.cs:606-609
      pv[n] += add;
      if (!last_row.empty()) last_row[n] = add;
    }
    end_period_pv = global_total / total_n;
    if (!last_row.empty()) {
      if (h->stream_panels)
        std::copy(last_row.begin(), last_row.end(), h->sinks[5] + (size_t)P * N);
      else
        CUDA_OK(cudaMemcpy(h->panels[5].ptr + (size_t)P * N, last_row.data(), N * sizeof(double), cudaMemcpyHostToDevice));
    }
    CUDA_OK(cudaMemcpy(h->pv.ptr, pv.data(), N * sizeof(double), cudaMemcpyHostToDevice));
  }
  {
    double sum = device_sum(h, h->pv.ptr, N, 0.0, 1);
    h->comm.all_reduce_host(&sum, 1, st);
    npv = sum / total_n;
    double ss = device_sum(h, h->pv.ptr, N, npv, 2);
    h->comm.all_reduce_host(&ss, 1, st);
    variance = total_n > 1 ? ss / (total_n - 1.0) : 0.0;
  }
  double back_sum = device_sum(h, v_first, N, 0.0, 1);
  h->comm.all_reduce_host(&back_sum, 1, st);
  double final_inv_sum = device_sum(h, h->final_inv.ptr, N, 0.0, 1);
  h->comm.all_reduce_host(&final_inv_sum, 1, st);

  h->scalars.snapped_inventories = snapped;
  h->scalars.ms_upload = h->sims[0].ms_upload + h->sims[1].ms_upload;
  h->scalars.npv = npv;
  h->scalars.std_error = std::sqrt(variance) / std::sqrt(total_n);
  h->scalars.backward_npv = back_sum / total_n;
  double *pr = &h->profile[(size_t)P * 5];  // LSV.cs:627-629
  pr[0] = final_inv_sum / total_n;
  pr[4] = end_period_pv;
  h->scalars.ms_pseudo_inverse = t_pinv.ms();
  h->scalars.ms_backward = t_back.ms() + h->scalars.ms_pseudo_inverse;
  h->scalars.ms_forward = t_fwd.ms();
  h->scalars.ms_device = t_dev.ms();
  {
    double total = 0.0;
    int64_t timed = 0;
    for (int p = 0; p < P; ++p) {
      if (!timed_step(p)) continue;
      float m = 0.f;
      if (cudaEventElapsedTime(&m, step_events[2 * p], step_events[2 * p + 1]) == cudaSuccess) total += m, ++timed;
    }
    h->scalars.ms_backward_kernel = total;
    h->scalars.backward_kernel_launches = timed;  // the launches that were timed (every launch, or a regular sample of them)
  }
  h->scalars.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
  h->has_results = true;
  if (progress) progress(1.0, user);  // LSV.cs:644
}

}  // namespace

// ================================================================================================ C ABI
extern "C" {

lsmc_status lsmc_create(const lsmc_problem *pb, int device, lsmc_handle **out) {
  if (!out) return LSMC_ERR_ARGUMENT;
  *out = nullptr;
  std::unique_ptr<lsmc_handle> h(new lsmc_handle);
  try {
    validate_problem(pb);
    h->model.reset(new StorageModel(pb->storage));
    h->N = pb->num_sims;
    h->P = pb->storage.num_periods;
    h->K = pb->num_basis;
    h->F = pb->num_factors;
    h->KP = padded_basis(h->K);
    h->acc_k = acc_width(h->K);
    h->current_active = pb->current_period_is_active != 0;
    h->discount_deltas = pb->discount_deltas != 0;
    h->compat_terminal = pb->reference_compat_terminal_pv != 0;
    h->flags = pb->simulation_data_returned;
    h->inventory = pb->inventory;
    h->tolerance = pb->numerical_tolerance;
    h->grid_spacing = pb->grid_spacing;
    h->extra = pb->extra_decisions;
    h->forward_price.assign(pb->forward_price, pb->forward_price + h->P + 1);
    h->settle_df.assign(pb->settle_df, pb->settle_df + h->P);
    h->powers.assign(pb->basis_powers, pb->basis_powers + (size_t)h->K * (1 + h->F));
    h->tables = build_tables(*h->model, *pb);  // inventory space, grids, decision sets: all validation happens here
    h->first_sim = h->current_active ? 1 : 0;
    h->sim_rows = h->P + 1 - h->first_sim;
    h->Gmax = h->tables.max_grid;
    h->D = h->tables.max_decisions;
    h->pitch = (h->N + kTilePaths - 1) / kTilePaths * kTilePaths;  // whole tiles: the TMA row copies never run off a row
    if ((uint64_t)h->Gmax * (uint64_t)h->pitch >= (uint64_t)1 << 32)
      throw Error(LSMC_ERR_ARGUMENT, "paths x inventory-grid points per device must stay below 2^32 (shard the paths over more GPUs)");
    h->device = device;

    require_device(device);
    CUDA_OK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CUDA_OK(cudaStreamCreateWithFlags(&h->aux_stream, cudaStreamNonBlocking));
    for (auto &e : h->ev_basis) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : h->ev_step) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : h->ev_window) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&h->ev_prepass, cudaEventDisableTiming));
    h->v_a.allocate((size_t)h->Gmax * h->pitch);
    h->v_b.allocate((size_t)h->Gmax * h->pitch);
    // the padding columns [N, pitch) of the value rows are read (never written) by the sweep kernels: keep them finite
    CUDA_OK(cudaMemset(h->v_a.ptr, 0, h->v_a.count * sizeof(double)));
    CUDA_OK(cudaMemset(h->v_b.ptr, 0, h->v_b.count * sizeof(double)));
    h->gamma.allocate((size_t)h->P * h->Gmax * h->K);
    h->rinv.allocate((size_t)h->P * h->K * h->K * 2);
    CUDA_OK(cudaMemset(h->rinv.ptr, 0, h->rinv.count * sizeof(double)));
    const int64_t num_tiles = (h->N + kSweepTilePaths - 1) / kSweepTilePaths;  // the finer of the two kernels' tilings
    h->partials.allocate((size_t)num_tiles * h->Gmax * h->acc_k);
    {  // test hook: LSMC_SWEEP_MMA=0 keeps the fused form on the DFMA row-ring kernel
      const char *env = std::getenv("LSMC_SWEEP_MMA");
      h->use_mma = h->D == 3 && !(h->flags & LSMC_REFERENCE_OPERATION_ORDER) && !(env && env[0] == '0');
    }
    h->point_tables.allocate(std::max((size_t)h->Gmax * point_table_bytes(h->K),
                                      (size_t)((h->Gmax + 7) / 8) * mma_block_record_bytes(h->K)) / sizeof(double));
    h->fold_counter.allocate(1);
    CUDA_OK(cudaMemset(h->fold_counter.ptr, 0, sizeof(unsigned int)));
    h->qbuf[0].allocate((size_t)h->K * h->pitch);
    h->qbuf[1].allocate((size_t)h->K * h->pitch);
    h->qbuf[2].allocate((size_t)h->K * h->pitch);
    h->pv.allocate(h->N);
    h->final_inv.allocate(h->N);
    h->error_flag.allocate(4);
    h->block_sums.allocate(512);
    const uint32_t panel_flags[6] = {LSMC_RETURN_INVENTORY, LSMC_RETURN_INJECT_WITHDRAW_VOLUME, LSMC_RETURN_CMDTY_CONSUMED,
                                     LSMC_RETURN_INVENTORY_LOSS, LSMC_RETURN_NET_VOLUME, LSMC_RETURN_PV};
    {
      int num_panels = (h->flags & LSMC_RETURN_DECISION_INDEX) ? 1 : 0;
      for (int i = 0; i < 6; ++i)
        if (h->flags & panel_flags[i]) ++num_panels;
      h->stream_panels = (h->flags & LSMC_STREAM_PANELS) != 0 && num_panels > 0;
      size_t free_bytes = 0, total_bytes = 0;
      CUDA_OK(cudaMemGetInfo(&free_bytes, &total_bytes));
      const size_t sims_bytes = 2 * (size_t)(1 + h->F) * h->sim_rows * h->N * sizeof(double);  // both simulation sets, allocated later
      size_t rows_per_panel = (size_t)h->P + 1;
      if (h->stream_panels) {
        // two ring halves of `stream_block` periods per panel, sized for about 2 GB of device memory in total
        const size_t budget = (size_t)2 << 30;
        h->stream_block = (int)std::max<size_t>(1, std::min<size_t>((size_t)h->P, budget / (2 * (size_t)num_panels * h->N * sizeof(double))));
        if (const char *env = std::getenv("LSMC_STREAM_BLOCK"))  // test hook: periods per streamed block
          h->stream_block = std::max(1, std::min(h->P, std::atoi(env)));
        rows_per_panel = 2 * (size_t)h->stream_block;
        CUDA_OK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
          CUDA_OK(cudaEventCreateWithFlags(&h->ev_block[i], cudaEventDisableTiming));
          CUDA_OK(cudaEventCreateWithFlags(&h->ev_copied[i], cudaEventDisableTiming));
        }
      }
      const size_t panel_bytes = (size_t)num_panels * rows_per_panel * h->N * sizeof(double);
      if (num_panels > 0 && panel_bytes + sims_bytes > free_bytes) {
        auto gb = [](size_t b) { return std::to_string((b + ((size_t)1 << 29)) >> 30) + " GB"; };
        throw Error(LSMC_ERR_ARGUMENT, "SimulationDataReturned asks for " + gb(panel_bytes) + " of per-simulation panels on the device on top of " +
                                           gb(sims_bytes) + " of simulations, with " + gb(free_bytes) + " free: request fewer panels, add "
                                           "LSMC_STREAM_PANELS (panels are then delivered to host arrays block by block), or shard the paths "
                                           "over more GPUs.");
      }
      for (int i = 0; i < 6; ++i)
        if (h->flags & panel_flags[i]) {
          h->panels[i].allocate(rows_per_panel * h->N);
          CUDA_OK(cudaMemset(h->panels[i].ptr, 0, h->panels[i].count * sizeof(double)));
        }
      if (h->flags & LSMC_RETURN_DECISION_INDEX) h->panels[6].allocate((h->stream_panels ? rows_per_panel : (size_t)h->P) * h->N);
    }
    build_device_tables(h.get());
    *out = h.release();
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

static lsmc_status set_sims_impl(lsmc_handle *h, int which, const double *spot, const double *const *factors, bool device_ptrs,
                                 bool async = false) {
  if (!h) return LSMC_ERR_ARGUMENT;
  try {
    if (which != 0 && which != 1) throw Error(LSMC_ERR_ARGUMENT, "which must be 0 (regression) or 1 (valuation)");
    if (!spot || (h->F > 0 && !factors)) throw Error(LSMC_ERR_ARGUMENT, which == 0 ? "regressionSpotSim is null" : "valuationSpotSim is null");
    CUDA_OK(cudaSetDevice(h->device));
    SimSet &s = h->sims[which];
    s.wait_upload();  // an earlier asynchronous upload into the same buffers
    const size_t count = (size_t)h->sim_rows * h->N;
    s.ms_upload = 0.0;
    if (device_ptrs) {
      s.spot = spot;
      for (int i = 0; i < h->F; ++i) s.factors[i] = factors[i];
    } else {
      for (int i = 0; i < h->F; ++i)
        if (!factors[i]) throw Error(LSMC_ERR_ARGUMENT, "a factor simulation panel is null");
      s.owned_spot.allocate(count);
      s.spot = s.owned_spot.ptr;
      for (int i = 0; i < h->F; ++i) {
        s.owned_factors[i].allocate(count);
        s.factors[i] = s.owned_factors[i].ptr;
      }
      const int F = h->F, device = h->device;
      std::vector<const double *> src(factors, factors + F);
      if (async) {
        if (!s.upload_stream) CUDA_OK(cudaStreamCreateWithFlags(&s.upload_stream, cudaStreamNonBlocking));
        if (!s.upload_stager) s.upload_stager.reset(new PinnedStager);
        SimSet *sp = &s;
        s.uploader = std::thread([sp, spot, src, count, F, device] {
          try {
            const auto t0 = std::chrono::steady_clock::now();
            CUDA_OK(cudaSetDevice(device));
            sp->upload_stager->to_device(sp->owned_spot.ptr, spot, count * sizeof(double), sp->upload_stream);
            for (int i = 0; i < F; ++i) sp->upload_stager->to_device(sp->owned_factors[i].ptr, src[i], count * sizeof(double), sp->upload_stream);
            sp->ms_upload = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
          } catch (...) {
            sp->upload_error = std::current_exception();
          }
        });
      } else {
        const auto t0 = std::chrono::steady_clock::now();
        h->stager.to_device(s.owned_spot.ptr, spot, count * sizeof(double), h->stream);
        for (int i = 0; i < F; ++i) h->stager.to_device(s.owned_factors[i].ptr, src[i], count * sizeof(double), h->stream);
        CUDA_OK(cudaStreamSynchronize(h->stream));
        s.ms_upload = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      }
    }
    s.ms_simulation = 0.0;
    s.ready = true;
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

lsmc_status lsmc_set_sims(lsmc_handle *h, int which, const double *spot, const double *const *factors) {
  return set_sims_impl(h, which, spot, factors, false);
}

lsmc_status lsmc_set_sims_device(lsmc_handle *h, int which, const double *spot, const double *const *factors) {
  return set_sims_impl(h, which, spot, factors, true);
}

lsmc_status lsmc_set_sims_async(lsmc_handle *h, int which, const double *spot, const double *const *factors) {
  return set_sims_impl(h, which, spot, factors, false, true);
}

lsmc_status lsmc_set_panel_sink(lsmc_handle *h, lsmc_panel which, double *host_destination) {
  if (!h) return LSMC_ERR_ARGUMENT;
  try {
    if (!h->stream_panels) throw Error(LSMC_ERR_INVALID_OPERATION, "lsmc_set_panel_sink needs LSMC_STREAM_PANELS in SimulationDataReturned.");
    int slot = -1;
    switch (which) {
      case LSMC_PANEL_INVENTORY: slot = 0; break;
      case LSMC_PANEL_INJECT_WITHDRAW_VOLUME: slot = 1; break;
      case LSMC_PANEL_CMDTY_CONSUMED: slot = 2; break;
      case LSMC_PANEL_INVENTORY_LOSS: slot = 3; break;
      case LSMC_PANEL_NET_VOLUME: slot = 4; break;
      case LSMC_PANEL_PV: slot = 5; break;
      case LSMC_PANEL_DECISION_INDEX: slot = 6; break;
      default: throw Error(LSMC_ERR_ARGUMENT, "only the per-simulation result panels can be streamed");
    }
    if (!h->panels[slot].ptr) throw Error(LSMC_ERR_INVALID_OPERATION, "This panel was not requested in SimulationDataReturned.");
    h->sinks[slot] = host_destination;
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

lsmc_status lsmc_simulate(lsmc_handle *h, int which, const lsmc_multi_factor_params *params, uint64_t seed, int antithetic,
                          int64_t first_path_id) {
  if (!h) return LSMC_ERR_ARGUMENT;
  try {
    if (which != 0 && which != 1) throw Error(LSMC_ERR_ARGUMENT, "which must be 0 (regression) or 1 (valuation)");
    if (!params) throw Error(LSMC_ERR_ARGUMENT, "modelParameters is null");
    if (params->num_factors != h->F) throw Error(LSMC_ERR_ARGUMENT, "model has a different number of factors than the problem");
    CUDA_OK(cudaSetDevice(h->device));
    SimSet &s = h->sims[which];
    s.wait_upload();
    s.ms_upload = 0.0;
    const size_t count = (size_t)h->sim_rows * h->N;
    s.owned_spot.allocate(count);
    double *fac[LSMC_MAX_FACTORS] = {};
    for (int i = 0; i < h->F; ++i) {
      s.owned_factors[i].allocate(count);
      fac[i] = s.owned_factors[i].ptr;
    }
    Timer t(h->stream);
    t.start();
    simulate_multi_factor(*params, h->P, h->first_sim, h->forward_price.data(), h->N, seed, antithetic != 0, first_path_id,
                          s.owned_spot.ptr, fac, h->stream);
    t.stop();
    CUDA_OK(cudaStreamSynchronize(h->stream));
    CUDA_OK(cudaGetLastError());
    s.ms_simulation = t.ms();
    s.spot = s.owned_spot.ptr;
    for (int i = 0; i < h->F; ++i) s.factors[i] = fac[i];
    s.ready = true;
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

lsmc_status lsmc_simulate_noalloc(int device, const lsmc_multi_factor_params *params, int32_t num_periods, int32_t current_period_is_active,
                                  const double *forward_price, int64_t num_sims, uint64_t seed, int antithetic, int64_t first_path_id,
                                  double *spot_device, double *const *factors_device, void *cuda_stream) {
  try {
    if (!params) throw Error(LSMC_ERR_ARGUMENT, "modelParameters is null");
    simulate_multi_factor_noalloc(device, *params, num_periods, current_period_is_active ? 1 : 0, forward_price, num_sims, seed,
                                  antithetic != 0, first_path_id, spot_device, factors_device, static_cast<cudaStream_t>(cuda_stream));
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

int32_t lsmc_simulate_noalloc_max_periods(void) { return simulate_noalloc_max_periods(); }

lsmc_status lsmc_run(lsmc_handle *h, lsmc_progress_fn progress, void *user, const volatile int *cancel) {
  if (!h) return LSMC_ERR_ARGUMENT;
  try {
    run_valuation(h, progress, user, cancel);
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

int32_t lsmc_num_sim_periods(const lsmc_handle *h) { return h ? h->sim_rows : 0; }
int32_t lsmc_num_periods(const lsmc_handle *h) { return h ? h->P : 0; }
int32_t lsmc_max_grid_points(const lsmc_handle *h) { return h ? h->Gmax : 0; }
int64_t lsmc_num_evaluations(const lsmc_handle *h) { return h ? h->tables.grid_point_steps * h->N : 0; }

int32_t lsmc_get_grid(const lsmc_handle *h, int32_t period, double *out, int32_t capacity) {
  if (!h || period < 0 || period > h->P) return 0;
  const std::vector<double> &g = h->tables.periods[period].grid;
  for (int i = 0; i < std::min<int>(capacity, (int)g.size()); ++i) out[i] = g[i];
  return (int32_t)g.size();
}

lsmc_status lsmc_get_results(lsmc_handle *h, lsmc_results *out) {
  if (!h || !out) return LSMC_ERR_ARGUMENT;
  try {
    if (!h->has_results) throw Error(LSMC_ERR_INVALID_OPERATION, "lsmc_run has not completed on this handle.");
    double *deltas = out->deltas, *profile = out->profile, *pv = out->pv_by_sim;
    lsmc_trigger_prices *trig = out->trigger_prices;
    *out = h->scalars;
    out->deltas = deltas;
    out->profile = profile;
    out->trigger_prices = trig;
    out->pv_by_sim = pv;
    if (deltas) std::copy(h->deltas.begin(), h->deltas.end(), deltas);
    if (profile) std::copy(h->profile.begin(), h->profile.end(), profile);
    if (trig) std::copy(h->trigger.begin(), h->trigger.end(), trig);
    if (pv) {
      CUDA_OK(cudaSetDevice(h->device));
      CUDA_OK(cudaMemcpy(pv, h->pv.ptr, h->N * sizeof(double), cudaMemcpyDeviceToHost));
    }
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

lsmc_status lsmc_copy_panel(lsmc_handle *h, lsmc_panel which, double *dst) {
  if (!h || !dst) return LSMC_ERR_ARGUMENT;
  try {
    if (!h->has_results) throw Error(LSMC_ERR_INVALID_OPERATION, "lsmc_run has not completed on this handle.");
    CUDA_OK(cudaSetDevice(h->device));
    const size_t N = h->N;
    if (h->stream_panels && which != LSMC_PANEL_SPOT_REGRESSION && which != LSMC_PANEL_SPOT_VALUATION &&
        which != LSMC_PANEL_FACTORS_REGRESSION && which != LSMC_PANEL_FACTORS_VALUATION)
      throw Error(LSMC_ERR_INVALID_OPERATION, "LSMC_STREAM_PANELS: this panel was delivered to its sink during lsmc_run and is not kept on the device.");
    auto copy = [&](const double *src, size_t count) {
      if (!src) throw Error(LSMC_ERR_INVALID_OPERATION, "This panel was not requested in SimulationDataReturned.");
      h->stager.to_host(dst, src, count * sizeof(double), h->stream);
    };
    switch (which) {
      case LSMC_PANEL_INVENTORY: copy(h->panels[0].ptr, (h->P + 1) * N); break;
      case LSMC_PANEL_INJECT_WITHDRAW_VOLUME: copy(h->panels[1].ptr, (h->P + 1) * N); break;
      case LSMC_PANEL_CMDTY_CONSUMED: copy(h->panels[2].ptr, (h->P + 1) * N); break;
      case LSMC_PANEL_INVENTORY_LOSS: copy(h->panels[3].ptr, (h->P + 1) * N); break;
      case LSMC_PANEL_NET_VOLUME: copy(h->panels[4].ptr, (h->P + 1) * N); break;
      case LSMC_PANEL_PV: copy(h->panels[5].ptr, (h->P + 1) * N); break;
      case LSMC_PANEL_DECISION_INDEX: copy(h->panels[6].ptr, (size_t)h->P * N); break;
      case LSMC_PANEL_SPOT_REGRESSION: copy(h->sims[0].spot, (size_t)h->sim_rows * N); break;
      case LSMC_PANEL_SPOT_VALUATION: copy(h->sims[1].spot, (size_t)h->sim_rows * N); break;
      case LSMC_PANEL_FACTORS_REGRESSION:
      case LSMC_PANEL_FACTORS_VALUATION: {
        const SimSet &s = h->sims[which == LSMC_PANEL_FACTORS_REGRESSION ? 0 : 1];
        for (int i = 0; i < h->F; ++i)
          h->stager.to_host(dst + (size_t)i * h->sim_rows * N, s.factors[i], (size_t)h->sim_rows * N * sizeof(double), h->stream);
        break;
      }
      default: throw Error(LSMC_ERR_ARGUMENT, "unknown panel");
    }
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

lsmc_status lsmc_get_regression(lsmc_handle *h, double *gamma, double *r_inverse) {
  if (!h) return LSMC_ERR_ARGUMENT;
  try {
    if (!h->has_results) throw Error(LSMC_ERR_INVALID_OPERATION, "lsmc_run has not completed on this handle.");
    if (gamma) std::copy(h->h_gamma.begin(), h->h_gamma.end(), gamma);
    if (r_inverse)
      for (size_t i = 0; i < h->h_rinv.size() / 2; ++i) r_inverse[i] = h->h_rinv[2 * i];
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

const char *lsmc_last_error(const lsmc_handle *h) { return h ? h->error.c_str() : g_create_error.c_str(); }

void lsmc_destroy(lsmc_handle *h) {
  if (!h) return;
  cudaSetDevice(h->device);
  for (auto &s : h->sims)  // an upload still in flight writes into buffers this call is about to free
    if (s.uploader.joinable()) s.uploader.join();
  if (h->stream) {
    cudaStreamSynchronize(h->stream);
    if (h->aux_stream) cudaStreamSynchronize(h->aux_stream);
    cudaStreamDestroy(h->stream);
    if (h->aux_stream) cudaStreamDestroy(h->aux_stream);
    for (auto &e : h->ev_basis) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_step) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_window) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_block) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_copied) if (e) cudaEventDestroy(e);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (auto &e : h->step_events) if (e) cudaEventDestroy(e);
    if (h->ev_prepass) cudaEventDestroy(h->ev_prepass);
  }
  delete h;
}

lsmc_status lsmc_comm_unique_id(void *id_out) {
  try {
    Comm::unique_id(id_out);
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

lsmc_status lsmc_comm_init(lsmc_handle *h, const void *id, int rank, int world, int64_t total_sims) {
  if (!h) return LSMC_ERR_ARGUMENT;
  try {
    CUDA_OK(cudaSetDevice(h->device));
    h->comm.init(id, rank, world, total_sims);
    h->comm.setup_peer_exchange((size_t)h->Gmax * h->K, h->stream);  // falls back to NCCL all-gathers when peers cannot be mapped
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(h, e);
  }
}

// ---- host helpers -----------------------------------------------------------------------------------------------
lsmc_status lsmc_calculate_inventory_space(const lsmc_storage *storage, double starting_inventory, double *out_min, double *out_max) {
  try {
    if (!storage) throw Error(LSMC_ERR_ARGUMENT, "storage is null");
    StorageModel st(*storage);
    std::vector<double> lo, hi;
    calculate_inventory_space(st, starting_inventory, lo, hi);
    std::copy(lo.begin(), lo.end(), out_min);
    std::copy(hi.begin(), hi.end(), out_max);
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

int32_t lsmc_bang_bang_decision_set(double min_rate, double max_rate, double inventory, double inventory_loss, double next_min,
                                    double next_max, double tolerance, int32_t extra, double *out) {
  try {
    double tmp[LSMC_MAX_DECISIONS];
    int n = bang_bang_decision_set({min_rate, max_rate}, inventory, inventory_loss, next_min, next_max, tolerance, extra, tmp);
    std::copy(tmp, tmp + n, out);
    return n;
  } catch (const std::exception &e) {
    return -status_of(nullptr, e);
  }
}

lsmc_status lsmc_bisect_inventory_space(const double *grid, int32_t n, double inventory, double tolerance, int32_t *lower, int32_t *upper) {
  try {
    int lo, hi;
    bisect_inventory_space(grid, n, inventory, tolerance, &lo, &hi);
    *lower = lo;
    *upper = hi;
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

int32_t lsmc_grid_points(double lower, double upper, double spacing, double *out, int32_t capacity) {
  try {
    std::vector<double> g = grid_points(lower, upper, spacing);
    for (int i = 0; i < std::min<int>(capacity, (int)g.size()); ++i) out[i] = g[i];
    return (int32_t)g.size();
  } catch (const std::exception &e) {
    return -status_of(nullptr, e);
  }
}

lsmc_status lsmc_inject_withdraw_range(const lsmc_storage *storage, int32_t period, double inventory, double *min_rate, double *max_rate) {
  try {
    if (!storage) throw Error(LSMC_ERR_ARGUMENT, "storage is null");
    StorageModel st(*storage);
    RateRange r = st.inject_withdraw_range(period, inventory);
    *min_rate = r.lo;
    *max_rate = r.hi;
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

lsmc_status lsmc_inventory_space_bounds(const lsmc_storage *storage, int32_t period, double next_lower, double next_upper,
                                        double *lower, double *upper) {
  try {
    if (!storage) throw Error(LSMC_ERR_ARGUMENT, "storage is null");
    StorageModel st(*storage);
    if (period < 0 || period >= st.num_periods()) throw Error(LSMC_ERR_ARGUMENT, "period out of range");
    *upper = st.space_upper_bound(period, next_lower, next_upper);
    *lower = st.space_lower_bound(period, next_lower, next_upper);
    return LSMC_OK;
  } catch (const std::exception &e) {
    return (lsmc_status)status_of(nullptr, e);
  }
}

int32_t lsmc_parse_basis(const char *expression, int32_t num_factors, int32_t *powers, int32_t capacity) {
  try {
    return parse_basis(expression, num_factors, powers, capacity);
  } catch (const std::exception &e) {
    return -status_of(nullptr, e);
  }
}

}  // extern "C"
