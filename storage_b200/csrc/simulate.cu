// simulate.cu — multi-factor spot price simulation on the device.
//
// Model (Cmdty.Core 0.2.0 MultiFactorSpotPriceSimulator, un-vendored; parameterisation documented in
// src/Cmdty.Storage.Python/cmdty_storage/multi_factor_diffusion_model.py:49-70):
//   dF(t,T)/F(t,T) = sum_i vol_i(T) exp(-a_i (T - t)) dz_i(t),   corr(dz_i, dz_j) = rho_ij
//   x_i(T) = int_0^T exp(-a_i (T - s)) dz_i(s)                   (Ornstein-Uhlenbeck Markov factors, simulated exactly)
//   S(T)   = F(0,T) exp( sum_i vol_i(T) x_i(T) - Var[sum_i vol_i(T) x_i(T)] / 2 )
// so that E[S(T)] = F(0,T) and Cov matches MultiFactorModel.integrated_covar.
//
// Random numbers: Philox4x32-10 keyed by the seed, counter = (antithetic pair id, period, draw) — a path's draws depend
// only on its global id, not on how paths are sharded over GPUs.  Antithetic: paths 2j and 2j+1 use z and -z.
// Layout: one thread per path, periods walked sequentially; every store is coalesced over paths (path-major FP64 rows).
#include "simulate.hpp"

#include <cmath>
#include <mutex>
#include <string>
#include <vector>

#include "host_model.hpp"

namespace lsmc {

namespace {

struct StepParams {  // per simulated period
  double decay[LSMC_MAX_FACTORS];                    // exp(-a_i dt)
  double chol[LSMC_MAX_FACTORS * LSMC_MAX_FACTORS];  // lower Cholesky factor of the increment covariance, row-major
  double vol[LSMC_MAX_FACTORS];                      // vol_i of the contract delivering in this period
  double forward;                                    // F(0,T)
  double half_var;                                   // Var / 2
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

__device__ __forceinline__ double uniform_open(uint32_t hi, uint32_t lo) {
  const uint64_t bits = ((uint64_t)hi << 32) | lo;
  return ((double)(bits >> 11) + 0.5) * (1.0 / 9007199254740992.0);  // (0,1), 53 bits
}

struct FactorOutputs {  // factor output rows, passed by value in the launch parameters
  double *ptr[LSMC_MAX_FACTORS];
};

// One path through all simulated periods: shared by simulate_kernel and simulate_noalloc_kernel, which differ only in where
// the per-period table lives.
__device__ __forceinline__ void simulate_path(const StepParams *__restrict__ steps, int rows, int F, int64_t N, uint2 key, int antithetic,
                                              int64_t first_path_id, int64_t n, double *__restrict__ spot, const FactorOutputs &out) {
  const uint64_t gid = (uint64_t)(first_path_id + n);
  const uint64_t stream_id = antithetic ? (gid >> 1) : gid;
  const double sign = (antithetic && (gid & 1)) ? -1.0 : 1.0;
  double x[LSMC_MAX_FACTORS];
#pragma unroll
  for (int i = 0; i < LSMC_MAX_FACTORS; ++i) x[i] = 0.0;
  for (int r = 0; r < rows; ++r) {
    const StepParams &sp = steps[r];
    double z[LSMC_MAX_FACTORS];
#pragma unroll
    for (int i = 0; i < LSMC_MAX_FACTORS; i += 2) {
      if (i < F) {
        const uint4 bits = philox4x32_10(make_uint4((uint32_t)stream_id, (uint32_t)(stream_id >> 32), (uint32_t)r, (uint32_t)(i >> 1)), key);
        const double u1 = uniform_open(bits.x, bits.y), u2 = uniform_open(bits.z, bits.w);
        const double radius = sqrt(-2.0 * log(u1));
        double sn, cs;
        sincospi(2.0 * u2, &sn, &cs);
        z[i] = sign * radius * cs;
        if (i + 1 < LSMC_MAX_FACTORS) z[i + 1] = sign * radius * sn;
      }
    }
    double log_shift = 0.0;
#pragma unroll
    for (int i = 0; i < LSMC_MAX_FACTORS; ++i) {
      if (i < F) {
        double inc = 0.0;
#pragma unroll
        for (int j = 0; j <= i; ++j) inc += sp.chol[i * LSMC_MAX_FACTORS + j] * z[j];
        x[i] = x[i] * sp.decay[i] + inc;
        log_shift += sp.vol[i] * x[i];
        out.ptr[i][(size_t)r * N + n] = x[i];
      }
    }
    spot[(size_t)r * N + n] = sp.forward * exp(log_shift - sp.half_var);
  }
}

__global__ void __launch_bounds__(256) simulate_kernel(const StepParams *__restrict__ steps, int rows, int F, int64_t N, uint2 key,
                                                       int antithetic, int64_t first_path_id, double *__restrict__ spot, FactorOutputs out) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n < N) simulate_path(steps, rows, F, N, key, antithetic, first_path_id, n, spot, out);
}

double integral_exp(double a, double t) {  // int_0^t exp(-a s) ds
  return a == 0.0 ? t : (1.0 - std::exp(-a * t)) / a;
}

// Parameters of simulated period p (exact OU transition over [t_prev, t], Cholesky factor of its covariance, drift correction);
// shared by both entry points.
void fill_step_params(const lsmc_multi_factor_params &mp, int P, int p, double t_prev, const double *forward_price, StepParams &sp) {
  const int F = mp.num_factors;
  const double t = mp.period_time[p];
  const double dt = t - t_prev;
  if (dt < 0.0) throw Error(LSMC_ERR_ARGUMENT, "period times must be non-decreasing and start on or after the current date");
  std::fill(std::begin(sp.decay), std::end(sp.decay), 0.0);
  std::fill(std::begin(sp.chol), std::end(sp.chol), 0.0);
  std::fill(std::begin(sp.vol), std::end(sp.vol), 0.0);
  double cov[LSMC_MAX_FACTORS][LSMC_MAX_FACTORS];
  for (int i = 0; i < F; ++i) {
    sp.decay[i] = std::exp(-mp.mean_reversion[i] * dt);
    sp.vol[i] = mp.factor_vol[(size_t)i * (P + 1) + p];
    for (int j = 0; j < F; ++j)
      cov[i][j] = mp.factor_corr[i * F + j] * integral_exp(mp.mean_reversion[i] + mp.mean_reversion[j], dt);
  }
  // Cholesky of the increment covariance (semi-definite tolerated: a zero pivot zeroes its column, e.g. dt == 0)
  for (int i = 0; i < F; ++i)
    for (int j = 0; j <= i; ++j) {
      double s = cov[i][j];
      for (int k = 0; k < j; ++k) s -= sp.chol[i * LSMC_MAX_FACTORS + k] * sp.chol[j * LSMC_MAX_FACTORS + k];
      if (i == j) {
        if (s < -1e-12 * std::fabs(cov[i][i])) throw Error(LSMC_ERR_ARGUMENT, "factor correlation matrix is not positive semi-definite");
        sp.chol[i * LSMC_MAX_FACTORS + i] = s > 0.0 ? std::sqrt(s) : 0.0;
      } else {
        const double d = sp.chol[j * LSMC_MAX_FACTORS + j];
        sp.chol[i * LSMC_MAX_FACTORS + j] = d > 0.0 ? s / d : 0.0;
      }
    }
  double var = 0.0;  // Var[sum_i vol_i x_i(t)] = sum_ij vol_i vol_j rho_ij (1 - exp(-(a_i+a_j) t)) / (a_i+a_j)
  for (int i = 0; i < F; ++i)
    for (int j = 0; j < F; ++j)
      var += sp.vol[i] * sp.vol[j] * mp.factor_corr[i * F + j] * integral_exp(mp.mean_reversion[i] + mp.mean_reversion[j], t);
  sp.half_var = 0.5 * var;
  sp.forward = forward_price[p];
}

}  // namespace

void simulate_multi_factor(const lsmc_multi_factor_params &mp, int P, int first_sim, const double *forward_price, int64_t N,
                           uint64_t seed, bool antithetic, int64_t first_path_id, double *spot, double *const *factors,
                           cudaStream_t stream) {
  const int F = mp.num_factors;
  if (F < 1 || F > LSMC_MAX_FACTORS) throw Error(LSMC_ERR_ARGUMENT, "number of factors must be in [1, LSMC_MAX_FACTORS]");
  if (!mp.mean_reversion || !mp.factor_vol || !mp.factor_corr || !mp.period_time) throw Error(LSMC_ERR_ARGUMENT, "model parameter arrays must not be null");
  const int rows = P + 1 - first_sim;
  std::vector<StepParams> steps(rows);
  double t_prev = 0.0;
  for (int r = 0; r < rows; ++r) {
    fill_step_params(mp, P, first_sim + r, t_prev, forward_price, steps[r]);
    t_prev = mp.period_time[first_sim + r];
  }
  StepParams *d_steps = nullptr;
  auto fail = [&](const char *what) {
    if (d_steps) cudaFree(d_steps);
    throw Error(LSMC_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(cudaGetLastError()));
  };
  if (cudaMalloc(&d_steps, sizeof(StepParams) * rows) != cudaSuccess) fail("cudaMalloc(step params)");
  FactorOutputs out{};
  for (int i = 0; i < F; ++i) out.ptr[i] = factors[i];
  if (cudaMemcpyAsync(d_steps, steps.data(), sizeof(StepParams) * rows, cudaMemcpyHostToDevice, stream) != cudaSuccess) fail("copy step params");
  const int threads = 256;
  const int blocks = (int)((N + threads - 1) / threads);
  simulate_kernel<<<blocks, threads, 0, stream>>>(d_steps, rows, F, N, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)),
                                                   antithetic ? 1 : 0, first_path_id, spot, out);
  if (cudaStreamSynchronize(stream) != cudaSuccess) fail("simulate_kernel");
  cudaFree(d_steps);
}


// ====================================================================================================================
// Allocation-free variant (lsmc_simulate_noalloc)
// ====================================================================================================================
// For hosts that must not allocate device memory once they are running: a pricing service that sized and reserved all of
// its GPU memory at start-up (cudaMalloc later can fail on a fragmented heap and it synchronises the whole device, which
// stalls every other valuation in flight), a host that handed the GPU heap to an arena it owns and hands out sub-buffers
// from, or a caller in a phase where the CUDA allocator may not be called at all.  lsmc_simulate() above needs a
// per-period parameter table next to its inputs and outputs and gets it from cudaMalloc; this variant keeps that table
// in a fixed array that is part of the library image.
//
// The array costs kNoAllocMaxRows * sizeof(StepParams) (about 1.3 MB) of device memory on every GPU that loads the
// library, from load until the process exits.  It is shared by all calls: each call overwrites the rows it needs (the
// rest keeps whatever an earlier call left there and is never read, so nothing has to be cleared), and two calls whose
// GPU work overlaps — different streams, or different host threads — would overwrite each other's table.  Serialise them.
constexpr int kNoAllocMaxRows = 2048;  // simulated periods per call (a 5-year daily valuation needs 1826)

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocMaxRows * sizeof(StepParams)];

// Same arithmetic as simulate_kernel (the shared simulate_path); the per-period table is the static scratch, whose device
// address is passed in, resolved once per device by the host code below.
__global__ void __launch_bounds__(256) simulate_noalloc_kernel(const StepParams *__restrict__ steps, int rows, int F, int64_t N,
                                                               uint2 key, int antithetic, int64_t first_path_id,
                                                               double *__restrict__ spot, FactorOutputs out) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n < N) simulate_path(steps, rows, F, N, key, antithetic, first_path_id, n, spot, out);
}

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
StepParams *g_noalloc_ptr[kMaxDevices] = {};  // device address of g_noalloc_scratch, per device

// In host code the name g_noalloc_scratch is only a handle for the runtime, not an address a kernel could use; the device
// address is looked up once per device, with that device current.
StepParams *noalloc_scratch_for(int device) {
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (device < 0 || device >= kMaxDevices) throw Error(LSMC_ERR_ARGUMENT, "device index out of range");
  if (!g_noalloc_ptr[device]) {
    void *ptr = nullptr;
    cudaError_t err = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (err != cudaSuccess || !ptr)
      throw Error(LSMC_ERR_CUDA, std::string("cudaGetSymbolAddress(scratch) failed: ") + cudaGetErrorString(err));
    g_noalloc_ptr[device] = static_cast<StepParams *>(ptr);
  }
  return g_noalloc_ptr[device];
}

}  // namespace

void simulate_multi_factor_noalloc(int device, const lsmc_multi_factor_params &mp, int P, int first_sim, const double *forward_price,
                                   int64_t N, uint64_t seed, bool antithetic, int64_t first_path_id, double *spot,
                                   double *const *factors, cudaStream_t stream) {
  const int F = mp.num_factors;
  if (F < 1 || F > LSMC_MAX_FACTORS) throw Error(LSMC_ERR_ARGUMENT, "number of factors must be in [1, LSMC_MAX_FACTORS]");
  if (!mp.mean_reversion || !mp.factor_vol || !mp.factor_corr || !mp.period_time || !forward_price || !spot || !factors)
    throw Error(LSMC_ERR_ARGUMENT, "model parameter arrays and output buffers must not be null");
  if (P < 1 || first_sim < 0 || first_sim > 1 || N <= 0) throw Error(LSMC_ERR_ARGUMENT, "invalid period or simulation count");
  const int rows = P + 1 - first_sim;
  if (rows > kNoAllocMaxRows)
    throw Error(LSMC_ERR_ARGUMENT, "lsmc_simulate_noalloc: " + std::to_string(rows) + " simulated periods exceed the fixed scratch of " +
                                       std::to_string(kNoAllocMaxRows) + " periods; use lsmc_simulate or split the horizon");
  struct DeviceGuard {  // the caller's current device is put back on every way out
    int previous = -1;
    DeviceGuard() { if (cudaGetDevice(&previous) != cudaSuccess) previous = -1; }
    ~DeviceGuard() { if (previous >= 0) cudaSetDevice(previous); }
  } guard;
  if (cudaSetDevice(device) != cudaSuccess) throw Error(LSMC_ERR_CUDA, "cudaSetDevice failed (no usable CUDA device)");
  StepParams *table = noalloc_scratch_for(device);

  // per-period parameters, computed on the host in small batches on the stack and copied into the scratch table in
  // stream order (a pageable source is consumed before cudaMemcpyAsync returns)
  constexpr int kBatch = 32;
  StepParams batch[kBatch];
  double t_prev = 0.0;
  for (int r0 = 0; r0 < rows; r0 += kBatch) {
    const int nb = std::min(kBatch, rows - r0);
    for (int b = 0; b < nb; ++b) {
      const int p = first_sim + r0 + b;
      fill_step_params(mp, P, p, t_prev, forward_price, batch[b]);
      t_prev = mp.period_time[p];
    }
    if (cudaMemcpyAsync(table + r0, batch, sizeof(StepParams) * nb, cudaMemcpyHostToDevice, stream) != cudaSuccess)
      throw Error(LSMC_ERR_CUDA, std::string("copy of the period table into the scratch failed: ") + cudaGetErrorString(cudaGetLastError()));
  }
  FactorOutputs out{};
  for (int i = 0; i < F; ++i) {
    if (!factors[i]) throw Error(LSMC_ERR_ARGUMENT, "a factor output buffer is null");
    out.ptr[i] = factors[i];
  }
  const int threads = 256;
  const int blocks = (int)((N + threads - 1) / threads);
  simulate_noalloc_kernel<<<blocks, threads, 0, stream>>>(table, rows, F, N, make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)),
                                                           antithetic ? 1 : 0, first_path_id, spot, out);
  if (cudaGetLastError() != cudaSuccess) throw Error(LSMC_ERR_CUDA, "launch of the simulation kernel failed");
  if (cudaStreamSynchronize(stream) != cudaSuccess)
    throw Error(LSMC_ERR_CUDA, std::string("simulation kernel failed: ") + cudaGetErrorString(cudaGetLastError()));
}

int simulate_noalloc_max_periods() { return kNoAllocMaxRows; }

}  // namespace lsmc
