// device_types.hpp — plain structs shared by the host orchestration (lsmc_api.cu) and the kernels (kernels.cu).
#pragma once
#include <cstdint>

#include "../../include/lsmc_b200.h"

namespace lsmc {

constexpr int kThreads = 128;            // threads per CTA in the path-parallel kernels
constexpr int kPathsPerThread = 2;       // double2 vectorisation over adjacent paths
constexpr int kTilePaths = kThreads * kPathsPerThread;
constexpr int kWarps = kThreads / 32;
constexpr int kAccK = 8;                 // regression partial sums are stored padded to 8 (or 16) per grid point
constexpr int kMaxGridTile = 16;         // inventory-grid points per CTA in the backward kernel
constexpr int kSweepTilePaths = 128;     // paths per CTA of the three-decision sweep kernel (backward_sweep.cu)

// Device copy of host_model.hpp's DecisionEntry (same layout).
struct DevDecision {
  double volume, cost_npv, consumed, lower_w, upper_w;
  int32_t lower, upper;
};

// How the orthonormalised basis row q[0..K) of one path is obtained for one period.
struct BasisSpec {
  const double *spot;                    // [N] spot prices of the period (NULL in mean mode)
  const double *factors[LSMC_MAX_FACTORS];  // [N] each
  const double *rinv;                    // [K*K*2] upper-triangular R^-1 as (hi,lo) pairs, row-major [l][j]
  int mean_mode;                         // 1: q = e_0 (current-period step: continuation value = sample mean)
};

// Window of v_next rows that a block of 8 grid points reads (all decisions), as the row-ring sweep kernel consumes it:
// hi is a running maximum, so block b+1 adds exactly the rows (hi(b), hi(b+1)]; the slots are the rows modulo the ring size.
struct SweepBlock {
  int32_t lo, hi, lo_slot, hi_slot;
};

struct BackwardStepArgs {
  int64_t N;            // paths on this device
  int64_t pitch;        // row pitch of the V and Q buffers, in doubles
  int G;                // grid points of this period
  int Gn;               // grid points of the next period (rows of v_next)
  int D;                // max decisions (row width of the decision table)
  int K;                // basis functions
  const double *v_next; // [Gn][pitch]
  double *v_this;       // [G][pitch]
  const double *gamma;  // [Gn][K] regression coefficients (orthonormal basis) of v_next on this period's basis
  const DevDecision *decisions;  // [G][D]
  const int32_t *num_decisions;  // [G]
  const double *inventory_cost;  // [G]
  const double *spot;   // [N] prices used for the immediate cash flow (NULL: use const_price)
  double const_price;   // forward price when the period is the current period
  double settle_df;
  const double *q_this; // [K][pitch] orthonormal basis rows of this period (NULL: mean mode, q = e_0)
  const double *q_prev; // [K][pitch] basis rows of period-1 for the regression sums (NULL with accumulate: mean mode)
  int accumulate;       // 0: no further regression (period 0)
  double *partials;     // [G][num_tiles][acc] per-tile sums of q_prev[k] * v_this[g] (num_tiles = gridDim.y)
  // FP64-mma sweep kernel: 128-path tiles of the step; the first whole_tiles get one CTA each, each of the others tail_split CTAs
  int64_t num_tiles, whole_tiles;
  int tail_split;
  int subtiles_per_cta; // inventory-grid sub-tiles one CTA sweeps (kMaxGridTile points each; groups of 8 in the ring and sweep kernels)
  const SweepBlock *sweep_blocks;  // row-ring sweep kernel: [blocks of 8 grid points] of this period
  const void *point_tables;  // sweep kernel: [G] per-grid-point records written by fold_tables_kernel
  int ring_rows;        // row-ring sweep kernel: rows in the shared-memory ring (0: direct-load kernels)
  // terminal mode: v_this[g][n] = TerminalStorageNpv(spot[n], grid_this[g]) instead of the decision step
  int terminal_mode;
  int fused_payoff;     // 1: fused-payoff arithmetic (backward_common.cuh point_fused); 0: the reference's operation order
  const double *grid_this;  // [G]
  double terminal[4];
};

// Per-period storage parameters for the forward pass (device-evaluable ICmdtyStorage).
struct DevPeriodParams {
  double min_inventory, max_inventory;
  double inj_cost, inj_df, wd_cost, wd_df, cons_inj, cons_wd, loss, inv_cost, inv_cost_df;
  double next_space_lo, next_space_hi;
  double settle_df, forward_price, delta_df;
  int32_t kind, rbegin, rcount, poly_begin;
  int32_t grid_next_count;   // points in the next period's grid
  int32_t mean_mode;         // period 0 when the current period is active
  int32_t sim_row;           // row of this period in the simulation panels (-1: not simulated)
  int32_t pad;
};

constexpr int kForwardSums = 8;  // sum_spot_vol, sum_vol, sum_cons, sum_loss, sum_pv, sum_inventory (+2 spare) then K basis sums

struct ForwardArgs {
  int64_t N;
  int P, K, F, Gmax;
  // Periods [p_begin, p_end) of the forward simulation in this launch.  A launch that does not start at period 0 resumes from
  // the per-path state (inventory, PV so far) the previous launch left in final_inventory / pv_by_sim; panel rows are written
  // relative to panel_row0 (streamed panels: the launch's first period; resident panels: 0).
  int p_begin, p_end, panel_row0;
  int64_t sim_pitch;               // N (caller layout)
  const DevPeriodParams *periods;  // [P]
  const double *ratchet_inventory, *ratchet_min, *ratchet_max;
  const double *ratchet_min_slope, *ratchet_max_slope;  // per ratchet segment: (y[k+1]-y[k]) / (x[k+1]-x[k]) (piecewise-linear kind)
  const double *poly_min, *poly_max;
  const double *grids;             // [P+1][Gmax]
  int first_pitch;                 // row pitch of bisect_first, a multiple of 16 (forward_first_pitch)
  int table_buffers;               // shared-memory copies of a period's tables: 2 = the next period's are fetched under this period's arithmetic
  const uint8_t *bisect_first;     // [P][first_pitch]: for the grid of period p+1, 1 where the reference's BisectInventorySpace meets mid = j
                                   //            before mid = j-1 (kernels.cu bracket_dev)
  const double *gamma;             // [P][Gmax][K]
  const double *rinv;              // [P][K*K*2]
  const int32_t *powers;           // [K][1+F]
  const double *spot;              // valuation sims [rows][N]
  const double *factors[LSMC_MAX_FACTORS];
  double start_inventory;
  double tolerance;
  double inv_grid_spacing;         // 1 / spacing of the fixed-spacing inventory grids
  int extra_decisions;
  double *pv_by_sim;               // [N]
  double *final_inventory;         // [N]
  double *step_sums;               // [P][num_ctas][kForwardSums + K]
  int sums_width;
  int tiles_per_cta;               // 128-path tiles each CTA carries through all periods (state kept in shared memory)
  // optional panels [P+1][N] (NULL = not requested)
  double *p_inventory, *p_volume, *p_consumed, *p_loss, *p_net_volume, *p_pv, *p_decision;
  int32_t *error_flag;             // set to an lsmc_status when a path leaves the feasible region
  unsigned long long *snapped;     // count of (path, period) pairs whose inventory was moved onto a limit
};

}  // namespace lsmc
