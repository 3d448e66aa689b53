// simulate.hpp — on-device multi-factor spot simulator (replaces Cmdty.Core's MultiFactorSpotPriceSimulator, call site
// LsmcValuationParameters.cs:164-183).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/lsmc_b200.h"

namespace lsmc {

// Writes spot[rows][N] and factors[i][rows][N] (rows = P+1-first_sim, row r = period first_sim+r) on `stream`.
void simulate_multi_factor(const lsmc_multi_factor_params &params, int P, int first_sim, const double *forward_price, int64_t N,
                           uint64_t seed, bool antithetic, int64_t first_path_id, double *spot, double *const *factors,
                           cudaStream_t stream);

// Allocation-free variant: same draws and arithmetic, outputs into caller-provided device buffers, the per-period
// parameter table in a fixed __device__ array of the library (at most simulate_noalloc_max_periods() simulated periods).
// Calls no cudaMalloc / cudaFree.  The first call on a device resolves the array's device address.
void simulate_multi_factor_noalloc(int device, const lsmc_multi_factor_params &params, int P, int first_sim, const double *forward_price,
                                   int64_t N, uint64_t seed, bool antithetic, int64_t first_path_id, double *spot,
                                   double *const *factors, cudaStream_t stream);
int simulate_noalloc_max_periods();

}  // namespace lsmc
