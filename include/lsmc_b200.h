/*
 * lsmc_b200.h — C ABI of liblsmc_b200.so: the B200-native Least-Squares Monte Carlo storage
 * valuation hot path (a drop-in for cmdty/storage's LsmcStorageValuation.Calculate).
 *
 * The reference has no native seam on this path: Calculate is one managed method.
 * This header therefore *defines* the seam.  Every entry point cites the reference interface it
 * replaces; citations are relative to /root/reference/src/Cmdty.Storage/ and use the shorthands
 *   LSV.cs = LsmcValuation/LsmcStorageValuation.cs      LVP.cs = LsmcValuation/LsmcValuationParameters.cs
 *   SH.cs  = StorageHelper.cs                           CS.cs  = StorageEntity/CmdtyStorage.cs
 *
 * Conventions
 *  - plain pointers and sizes only; all arrays are caller-owned and only read during the call that
 *    receives them (the library copies what it keeps); results are copied out into caller buffers.
 *  - every function returns an lsmc_status; lsmc_last_error() gives the message a managed host
 *    would put in the exception it maps the status to.
 *  - periods are indexed from the first *active* period: index 0 = max(storage.StartPeriod,
 *    CurrentPeriod) (SH.cs:45), index P = storage.EndPeriod; decisions are taken in 0..P-1.
 *  - panels are row-major [period][sim], sims contiguous — the layout of the reference's
 *    Panel<T,double> (PythonHelpers/SpotSimResultsFromPanels.cs:101-104).
 *  - a handle is used by one host thread at a time; distinct handles are independent.
 */
#ifndef LSMC_B200_H
#define LSMC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LSMC_MAX_BASIS 16      /* K: basis functions (monomials) per regression            */
#define LSMC_MAX_FACTORS 8     /* F: Markov factors                                         */
#define LSMC_MAX_DECISIONS 16  /* D: size of a bang-bang decision set, 2*ExtraDecisions + 3 */
#define LSMC_NUM_TRIGGER_VOLUMES 10 /* LSV.cs:383 numTriggerPriceVolumes */

/* Status codes and the reference exception each maps to. */
typedef enum lsmc_status {
  LSMC_OK = 0,
  LSMC_ERR_ARGUMENT = 1,              /* ArgumentException: LSV.cs:65,95,98; SH.cs:113,131,155,309,313; CS.cs:90,94 */
  LSMC_ERR_INVENTORY_CONSTRAINTS = 2, /* InventoryConstraintsCannotBeFulfilledException: LSV.cs:78; SH.cs:102       */
  LSMC_ERR_INVALID_OPERATION = 3,     /* InvalidOperationException from Builder.Build(): LVP.cs:126-138             */
  LSMC_ERR_CANCELLED = 4,             /* OperationCanceledException: LSV.cs:345,521                                 */
  LSMC_ERR_CUDA = 5,                  /* no reference counterpart: CUDA runtime failure / no device                 */
  LSMC_ERR_NUMERICAL = 6,             /* rank-deficient design matrix (the reference's QR would return Inf/NaN)     */
  LSMC_ERR_UNSUPPORTED = 7            /* opaque Func<> storage overloads that have no structured form (CS.cs:371,411,473,539) */
} lsmc_status;

/* IInjectWithdrawConstraint implementations (the four classes under InjectWithdrawConstraints/). */
typedef enum lsmc_constraint_kind {
  LSMC_CONSTRAINT_CONSTANT = 0,         /* ConstantInjectWithdrawConstraint.cs:45-65         */
  LSMC_CONSTRAINT_PIECEWISE_LINEAR = 1, /* PiecewiseLinearInjectWithdrawConstraint.cs:67-160 */
  LSMC_CONSTRAINT_STEP = 2,             /* StepInjectWithdrawConstraint.cs:72-166            */
  LSMC_CONSTRAINT_POLYNOMIAL = 3        /* PolynomialInjectWithdrawConstraint.cs:80-155      */
} lsmc_constraint_kind;

/* SimulationDataReturned flag values (LsmcValuation/SimulationDataReturned.cs:33-46). */
enum {
  LSMC_RETURN_NONE = 0,
  LSMC_RETURN_SPOT_REGRESSION = 1,
  LSMC_RETURN_SPOT_VALUATION = 1 << 2,
  LSMC_RETURN_FACTORS_REGRESSION = 1 << 3,
  LSMC_RETURN_FACTORS_VALUATION = 1 << 4,
  LSMC_RETURN_INVENTORY = 1 << 5,
  LSMC_RETURN_INJECT_WITHDRAW_VOLUME = 1 << 6,
  LSMC_RETURN_CMDTY_CONSUMED = 1 << 7,
  LSMC_RETURN_INVENTORY_LOSS = 1 << 8,
  LSMC_RETURN_NET_VOLUME = 1 << 9,
  LSMC_RETURN_PV = 1 << 10,
  LSMC_RETURN_DECISION_INDEX = 1 << 16, /* no reference counterpart: keep the argmax index of every (period, sim) */
  LSMC_REFERENCE_OPERATION_ORDER = 1 << 17, /* no reference counterpart: evaluate each decision's immediate cash flow with the
                                               reference's own sequence of roundings (LSV.cs:317-336) instead of the fused
                                               form S*coef + offset; both are within 1e-9 of the reference, the fused one is
                                               the default because it needs a third fewer FP64 operations */
  LSMC_STREAM_PANELS = 1 << 18, /* no reference counterpart: do not keep the requested per-simulation result panels on the device
                                   (6 x (P+1) x N doubles: 88 GB at 1M paths x 1825 periods); run the forward simulation in blocks
                                   of periods and deliver every finished block to the host arrays registered with
                                   lsmc_set_panel_sink while the next block is computed */
  LSMC_RETURN_ALL = 1 | (1 << 2) | (1 << 3) | (1 << 4) | (1 << 5) | (1 << 6) | (1 << 7) | (1 << 8) | (1 << 9) | (1 << 10)
};

/*
 * Structured form of ICmdtyStorage<T> (StorageEntity/ICmdtyStorage.cs:32-49) as built by
 * CmdtyStorage<T>.Builder (CS.cs:171-569).  The reference stores opaque Func<>s; the forward pass
 * evaluates the storage at off-grid per-path inventories (LSV.cs:450-471) so the device needs the
 * parameters, not the delegates.  Builder overloads taking arbitrary Func<>s (CS.cs:371,411,418,
 * 473,504,539) have no structured form and must be rejected host-side with LSMC_ERR_UNSUPPORTED.
 */
typedef struct lsmc_storage {
  int32_t num_periods;          /* P = EndPeriod.OffsetFrom(startActive) (SH.cs:47)                          */
  int32_t must_be_empty_at_end; /* CS.cs:424-428; also forces MaxInventory(end)=0 (CS.cs:435-438)            */
  const double *min_inventory;  /* [P+1] MinInventory(period) (CS.cs:107)                                   */
  const double *max_inventory;  /* [P+1] builder value; the MustBeEmptyAtEnd override is applied internally */
  /* Inject/withdraw constraint of period p in 0..P-1 = rows [ratchet_begin[p], ratchet_begin[p]+ratchet_count[p])
     of the ratchet table; periods may share rows (CmdtyStorageBuilderExtensions.cs:142-257).
     Rows of one constraint must be sorted by inventory (PiecewiseLinear…cs:45, Step…cs:42).
     CONSTANT uses the first row's rates only.  POLYNOMIAL rows are the points the polynomial is
     fitted through (Polynomial…cs:62-73). */
  const int32_t *constraint_kind; /* [P] lsmc_constraint_kind */
  const int32_t *ratchet_begin;   /* [P] */
  const int32_t *ratchet_count;   /* [P] */
  int32_t num_ratchet_rows;
  const double *ratchet_inventory; /* [num_ratchet_rows] */
  const double *ratchet_min_rate;  /* [num_ratchet_rows] MinInjectWithdrawRate (<=0: max withdrawal) */
  const double *ratchet_max_rate;  /* [num_ratchet_rows] MaxInjectWithdrawRate (>=0: max injection)  */
  /* Costs: the reference returns cash flows (date, amount); their NPV is amount * DF(current -> date)
     (LSV.cs:765-774).  The host resolves dates, so each cost is a per-unit amount and a discount factor. */
  const double *injection_cost;      /* [P] per unit injected (CS.cs:322-369)                    */
  const double *injection_cost_df;   /* [P] */
  const double *withdrawal_cost;     /* [P] per unit withdrawn (CS.cs:378-409)                   */
  const double *withdrawal_cost_df;  /* [P] */
  const double *consumed_on_inject;  /* [P] fraction of injected volume consumed (CS.cs:450-471) */
  const double *consumed_on_withdraw;/* [P] fraction of withdrawn volume consumed (CS.cs:480-502)*/
  const double *inventory_loss;      /* [P] fraction of inventory lost per period (CS.cs:511-537)*/
  const double *inventory_cost;      /* [P] per unit of inventory held (CS.cs:546-568)           */
  const double *inventory_cost_df;   /* [P] */
  /* TerminalStorageNpv(price, inv) = ((c[3]*(price*inv) + c[1]*price) + c[2]*inv) + c[0]
     (CS.cs:154-157,418-432; default all zero).  Covers the reference's own test lambda
     price*inv - 999 (tests/…/LsmcStorageValuationTest.cs:125). */
  double terminal_npv_coeff[4];
  /* PolynomialInjectWithdrawConstraint root finder parameters (Polynomial…cs:45-46). */
  double newton_raphson_accuracy;
  int32_t newton_raphson_max_iterations;
  int32_t newton_raphson_subdivision;
} lsmc_storage;

/*
 * Structured form of LsmcValuationParameters<T> (LVP.cs:38-80).  Date arithmetic (SettleDateRule,
 * DiscountFactors) is resolved by the host into per-period numbers.
 */
typedef struct lsmc_problem {
  lsmc_storage storage;
  double inventory;                 /* LVP.cs:43 starting inventory                                         */
  int32_t current_period_is_active; /* 1 if CurrentPeriod >= storage.StartPeriod: period 0 is not simulated,
                                       its price is ForwardCurve[current] and its continuation value is the
                                       sample mean (LSV.cs:174-185,232-236,397-405,424-428)                  */
  const double *forward_price;      /* [P+1] ForwardCurve[period] (LSV.cs:234,513)                          */
  const double *settle_df;          /* [P] DF(current -> SettleDateRule(period)) (LSV.cs:228-229,418-419)   */
  double grid_spacing;              /* FixedSpacingStateSpaceGridCalc.Spacing (StateSpaceGridCalc/Fixed…cs:36) */
  double numerical_tolerance;       /* LVP.cs:49,88 (default 1e-10)                                         */
  int32_t extra_decisions;          /* LVP.cs:56                                                            */
  int32_t discount_deltas;          /* LVP.cs:55                                                            */
  uint32_t simulation_data_returned;/* LVP.cs:57, LSMC_RETURN_* flags                                       */
  /* Basis functions (LVP.cs:52): monomial k = spot^basis_powers[k*(1+F)] * prod_i x_i^basis_powers[k*(1+F)+1+i]
     — the (SpotPower, MarkovFactorPowers) of PowerMonomialBuilder (BasisFunctions/PowerMonomialBuilder.cs:33-34),
     evaluated as BasisFunctions.Generic does (BasisFunctions/BasisFunctions.cs:93-135). */
  int32_t num_factors;
  int32_t num_basis;
  const int32_t *basis_powers;      /* [num_basis][1+num_factors] */
  int64_t num_sims;                 /* paths owned by this handle (this rank's shard when sharded)          */
  int32_t reference_compat_terminal_pv; /* 1: reproduce the cumulative terminal-PV sum of LSV.cs:606-609;
                                           0: each path receives its own terminal NPV                        */
} lsmc_problem;

/* Parameters of the multi-factor spot model simulated by Cmdty.Core 0.2.0's
 * MultiFactorSpotPriceSimulator<T> (un-vendored; call site LVP.cs:164-183):
 *   dF(t,T)/F(t,T) = sum_i vol_i(T) exp(-mr_i (T-t)) dz_i(t),  corr(dz_i,dz_j) = rho_ij,
 *   S(T) = F(0,T) exp( sum_i vol_i(T) x_i(T) - Var/2 ),  x_i(T) = int_0^T exp(-mr_i (T-s)) dz_i(s). */
typedef struct lsmc_multi_factor_params {
  int32_t num_factors;
  const double *mean_reversion; /* [F]                                                                  */
  const double *factor_vol;     /* [F][P+1] vol_i of the contract delivering in period p                */
  const double *factor_corr;    /* [F][F]                                                               */
  const double *period_time;    /* [P+1] Act/365 year fraction from CurrentPeriod.Start to period start (LVP.cs:176,180) */
} lsmc_multi_factor_params;

/* One TriggerPricePoint list + the TriggerPrices summary of a period (TriggerPrices.cs, LSV.cs:523-592). */
typedef struct lsmc_trigger_prices {
  int32_t num_inject;    /* 0 or LSMC_NUM_TRIGGER_VOLUMES */
  int32_t num_withdraw;  /* 0 or LSMC_NUM_TRIGGER_VOLUMES */
  double inject_volume[LSMC_NUM_TRIGGER_VOLUMES];
  double inject_price[LSMC_NUM_TRIGGER_VOLUMES];
  double withdraw_volume[LSMC_NUM_TRIGGER_VOLUMES];
  double withdraw_price[LSMC_NUM_TRIGGER_VOLUMES];
  int32_t has_inject_price;   /* TriggerPrices.HasInjectPrice   */
  int32_t has_withdraw_price; /* TriggerPrices.HasWithdrawPrice */
  double max_inject_volume, max_inject_trigger_price;
  double max_withdraw_volume, max_withdraw_trigger_price;
} lsmc_trigger_prices;

/* Scalar part of LsmcStorageValuationResults<T> (LsmcValuation/LsmcStorageValuationResults.cs:37-53).
 * Array members are caller-allocated; a NULL member is skipped. */
typedef struct lsmc_results {
  double npv;            /* Npv: forward-simulation mean (LSV.cs:617)                                  */
  double std_error;      /* ValuationSimStandardError: sample stdev / sqrt(N) (LSV.cs:618)             */
  double backward_npv;   /* the value the reference only logs (LSV.cs:623-625)                         */
  double *deltas;        /* [P+1] Deltas (LSV.cs:370,517-518); entry P is 0                            */
  double *profile;       /* [P+1][5] ExpectedStorageProfile: inventory, inject_withdraw_volume,
                            cmdty_consumed, inventory_loss, period_pv (LSV.cs:510-512,627-629)         */
  lsmc_trigger_prices *trigger_prices; /* [P] TriggerPrices + TriggerPriceVolumeProfiles (LSV.cs:523-592) */
  double *pv_by_sim;     /* [num_sims] PvBySim (LSV.cs:368,507,609)                                    */
  /* Stopwatches.cs:35-80 buckets, milliseconds (device time for GPU phases). */
  double ms_total, ms_regression_sim, ms_valuation_sim, ms_backward, ms_pseudo_inverse, ms_forward;
  int64_t kernel_launches; /* kernels of this library launched by the last lsmc_run */
  /* Device time (CUDA events on the launching stream) of the dominant kernel, the fused backward-induction step, summed
     over its TIMED launches of the last run, and the number of those launches (every launch up to 64 periods, every 8th
     beyond; LSMC_STEP_EVENTS=all times all of them); ms_device is the event time from the first to the last kernel of the run. */
  double ms_backward_kernel;
  int64_t backward_kernel_launches;
  double ms_device;
  /* Departure counter (DESIGN.md §2 "rounding at the inventory limits"): (path, period) pairs of the forward simulation whose
     inventory had left [MinInventory, MaxInventory] by less than NumericalTolerance and was moved onto the limit, where the
     reference throws (CmdtyStorage.cs:88-94).  0 means the run never relied on the departure. */
  int64_t snapped_inventories;
  double ms_upload; /* host wall time spent staging caller-supplied simulation panels to the device (lsmc_set_sims, lsmc_set_sims_async) */
} lsmc_results;

/* Per-sim panels selectable with SimulationDataReturned (LSV.cs:356-366,636-643). */
typedef enum lsmc_panel {
  LSMC_PANEL_INVENTORY = 0,              /* InventoryBySim             [P+1][N] */
  LSMC_PANEL_INJECT_WITHDRAW_VOLUME = 1, /* InjectWithdrawVolumeBySim  [P+1][N] (row P zero) */
  LSMC_PANEL_CMDTY_CONSUMED = 2,         /* CmdtyConsumedBySim         [P+1][N] */
  LSMC_PANEL_INVENTORY_LOSS = 3,         /* InventoryLossBySim         [P+1][N] */
  LSMC_PANEL_NET_VOLUME = 4,             /* NetVolumeBySim             [P+1][N] */
  LSMC_PANEL_PV = 5,                     /* PvByPeriodAndSim           [P+1][N] */
  LSMC_PANEL_SPOT_REGRESSION = 6,        /* RegressionSpotPriceSim     [rows][N], rows = lsmc_num_sim_periods */
  LSMC_PANEL_SPOT_VALUATION = 7,         /* ValuationSpotPriceSim      [rows][N] */
  LSMC_PANEL_FACTORS_REGRESSION = 8,     /* RegressionMarkovFactors    [F][rows][N] */
  LSMC_PANEL_FACTORS_VALUATION = 9,      /* ValuationMarkovFactors     [F][rows][N] */
  LSMC_PANEL_DECISION_INDEX = 10         /* no reference counterpart: argmax index per (period, sim) as
                                            doubles [P][N], for the bit-exact decision parity check */
} lsmc_panel;

typedef struct lsmc_handle lsmc_handle;

/* OnProgressUpdate (LVP.cs:54; LSV.cs:343-344,519-520,644): 2P+1 calls, last one 1.0. */
typedef void (*lsmc_progress_fn)(double progress, void *user);

/* ---- valuation -------------------------------------------------------------------------------- */

/* Replaces the set-up half of LsmcStorageValuation.Calculate (LSV.cs:57-98,215-306): validates,
 * computes the inventory space (SH.cs:39-107), the per-period grids (Fixed…GridCalc.cs:45-63) and the
 * sim-independent decision tables, and allocates the device buffers on CUDA device `device`. */
lsmc_status lsmc_create(const lsmc_problem *problem, int device, lsmc_handle **out);

/* Replaces LsmcValuationParameters.Builder.UseSpotSimResults (LVP.cs:198-216): caller-supplied
 * simulations.  which = 0 regression sims, 1 valuation sims.  spot is [rows][N], factors[i] is
 * [rows][N], rows = lsmc_num_sim_periods(h) covering max(current+1,start)..end (LVP.cs:219-231).
 * Pointers are host memory; the data is copied to the device inside the call. */
lsmc_status lsmc_set_sims(lsmc_handle *h, int which, const double *spot, const double *const *factors);

/* Same, for simulations already resident in device memory (device pointers, same layout).  The
 * handle keeps the pointers: the buffers must stay valid until lsmc_run returns. */
lsmc_status lsmc_set_sims_device(lsmc_handle *h, int which, const double *spot, const double *const *factors);

/* lsmc_set_sims without the wait (the caller-supplied-simulation overload at production size moves 58 GB per set):
 * returns at once, the panels are staged to the device by a worker of the handle.  The arrays must stay valid and unchanged
 * until lsmc_run has returned.  lsmc_run waits for the regression set before its first kernel and for the valuation set before
 * the forward simulation, so a valuation set given this way is uploaded while the backward induction runs. */
lsmc_status lsmc_set_sims_async(lsmc_handle *h, int which, const double *spot, const double *const *factors);

/* With LSMC_STREAM_PANELS (LsmcStorageValuation.cs:356-377,492-507,636-643 at sizes whose panels do not fit on the device):
 * where lsmc_run delivers panel `which` (LSMC_PANEL_INVENTORY .. LSMC_PANEL_PV, LSMC_PANEL_DECISION_INDEX), a host array of the
 * shape lsmc_copy_panel documents.  Must be called for every requested panel before lsmc_run; the array must stay valid until
 * lsmc_run has returned.  lsmc_copy_panel is not available for streamed panels. */
lsmc_status lsmc_set_panel_sink(lsmc_handle *h, lsmc_panel which, double *host_destination);

/* Replaces SimulateWithMultiFactorModelAndMersenneTwister (LVP.cs:185-196,164-183): simulate on the
 * device.  Counter-based RNG keyed by (seed, global path id, period) so a path's draws do not depend
 * on how paths are sharded; first_path_id is this shard's offset.  antithetic pairs paths (2j, 2j+1). */
lsmc_status lsmc_simulate(lsmc_handle *h, int which, const lsmc_multi_factor_params *params,
                          uint64_t seed, int antithetic, int64_t first_path_id);

/* No counterpart in the reference.  The same simulation as lsmc_simulate (same draws for the same seed and path ids),
 * for hosts that may not allocate device memory while they run: services that reserve all their GPU memory at start-up
 * because a later cudaMalloc can fail on a fragmented heap and synchronises the whole device, and hosts that own the
 * GPU heap through an arena of their own.  Nothing is allocated or freed on the device: the caller passes the output
 * buffers (device pointers: spot [rows][num_sims] and one [rows][num_sims] buffer per factor, rows = num_periods + 1 -
 * (current_period_is_active ? 1 : 0)), and the per-period parameter table the kernel needs lives in a fixed array inside
 * the library.  That array holds at most lsmc_simulate_noalloc_max_periods() (2048) simulated periods — a longer horizon
 * is rejected with LSMC_ERR_ARGUMENT — and occupies about 1.3 MB on every GPU that has the library loaded, until exit.
 * All calls share it and it keeps the previous call's rows (each call rewrites every row it reads, so nothing needs
 * clearing): calls whose GPU work overlaps, on different streams or from different host threads, corrupt each other's
 * table and must be serialised by the caller.  The first call on a device looks up the array's device address (once per
 * device, kept for later calls).  The kernel runs on `cuda_stream` (a cudaStream_t, NULL = default
 * stream); the call waits for it (it synchronises that stream) and returns after it has finished.  `device` is made current
 * for the duration of the call and the caller's current device is put back before it returns.  Errors: message via
 * lsmc_last_error(NULL). */
lsmc_status lsmc_simulate_noalloc(int device, const lsmc_multi_factor_params *params, int32_t num_periods,
                                  int32_t current_period_is_active, const double *forward_price, int64_t num_sims, uint64_t seed,
                                  int antithetic, int64_t first_path_id, double *spot_device, double *const *factors_device,
                                  void *cuda_stream);
int32_t lsmc_simulate_noalloc_max_periods(void);

/* Replaces the compute half of Calculate (LSV.cs:100-657): backward induction, forward simulation,
 * trigger prices, NPV.  `cancel` (may be NULL) is polled once per backward and forward step like
 * CancellationToken.ThrowIfCancellationRequested (LSV.cs:345,521). */
lsmc_status lsmc_run(lsmc_handle *h, lsmc_progress_fn progress, void *user, const volatile int *cancel);

/* Number of simulated periods (rows of a sim panel): P+1 if the current period is before the
 * storage start, else P (LVP.cs:177-178). */
int32_t lsmc_num_sim_periods(const lsmc_handle *h);
int32_t lsmc_num_periods(const lsmc_handle *h);
/* Inventory grid of period p (LSV.cs:215-223): returns the count, copies min(count,capacity) points. */
int32_t lsmc_get_grid(const lsmc_handle *h, int32_t period, double *out, int32_t capacity);
/* Total path x step x inventory-point evaluations of one run (the BASELINE.json metric's numerator). */
int64_t lsmc_num_evaluations(const lsmc_handle *h);

/* LsmcStorageValuationResults<T> (LsmcStorageValuationResults.cs:37-53). */
lsmc_status lsmc_get_results(lsmc_handle *h, lsmc_results *out);
/* Copies a per-sim panel to host memory `dst` (sized as documented at lsmc_panel). */
lsmc_status lsmc_copy_panel(lsmc_handle *h, lsmc_panel which, double *dst);
/* Regression coefficients in the orthonormalised basis and the matrix mapping monomials to it:
 * gamma [P][max_grid][K] and r_inverse [P][K][K]; the reference keeps raw-basis coefficients
 * (LSV.cs:199-212) but never returns them. max_grid = lsmc_max_grid_points. */
int32_t lsmc_max_grid_points(const lsmc_handle *h);
lsmc_status lsmc_get_regression(lsmc_handle *h, double *gamma, double *r_inverse);

const char *lsmc_last_error(const lsmc_handle *h); /* h == NULL: error of the last failed lsmc_create on this thread */
void lsmc_destroy(lsmc_handle *h);

/* ---- multi-GPU: one process per GPU, paths sharded ---------------------------- */

/* Size of the opaque id rank 0 obtains with lsmc_comm_unique_id and the host broadcasts
 * (torch.distributed / MPI / files) to the other ranks. */
#define LSMC_COMM_ID_BYTES 128
lsmc_status lsmc_comm_unique_id(void *id_out);
/* Joins this handle to a communicator of `world` ranks; total_sims is the path count over all
 * ranks (the N of every mean).  After this, lsmc_run exchanges the per-step regression partials
 * and the final sums across ranks. */
lsmc_status lsmc_comm_init(lsmc_handle *h, const void *id, int rank, int world, int64_t total_sims);

/* ---- host helpers: the public static surface of StorageHelper and friends ---------------------- */
/* These are the library's own host implementations used by lsmc_create; exported so a host can
 * expose the reference's StorageHelper API without re-implementing it, and so they can be tested
 * against the reference's unit-test vectors without a GPU. */

/* StorageHelper.CalculateInventorySpace (SH.cs:39-107): out_min/out_max [P] for periods 1..P. */
lsmc_status lsmc_calculate_inventory_space(const lsmc_storage *storage, double starting_inventory,
                                           double *out_min, double *out_max);
/* StorageHelper.CalculateBangBangDecisionSet (SH.cs:109-197): returns the count (<0: -lsmc_status). */
int32_t lsmc_bang_bang_decision_set(double min_rate, double max_rate, double inventory, double inventory_loss,
                                    double next_min_inventory, double next_max_inventory,
                                    double numerical_tolerance, int32_t extra_decisions, double *out);
/* StorageHelper.BisectInventorySpace (SH.cs:280-314). */
lsmc_status lsmc_bisect_inventory_space(const double *grid, int32_t n, double inventory, double tolerance,
                                        int32_t *lower, int32_t *upper);
/* FixedSpacingStateSpaceGridCalc.GetGridPoints (Fixed…GridCalc.cs:45-63): returns the count. */
int32_t lsmc_grid_points(double lower, double upper, double spacing, double *out, int32_t capacity);
/* IInjectWithdrawConstraint.GetInjectWithdrawRange of period p at `inventory` (CS.cs:86-100). */
lsmc_status lsmc_inject_withdraw_range(const lsmc_storage *storage, int32_t period, double inventory,
                                       double *min_rate, double *max_rate);
/* ICmdtyStorage.InventorySpaceLowerBound / InventorySpaceUpperBound of period p (CS.cs:132-152), which dispatch to the
 * period's constraint (Constant…cs:50-65, PiecewiseLinear…cs:74-160, Step…cs:81-166, Polynomial…cs:87-155). */
lsmc_status lsmc_inventory_space_bounds(const lsmc_storage *storage, int32_t period, double next_lower, double next_upper,
                                        double *lower, double *upper);
/* BasisFunctionsBuilder.Parse (BasisFunctions/BasisFunctionsBuilder.cs:90-131) without Roslyn:
 * "1 + s + x0**2 + x0*x1" -> exponent table [K][1+num_factors]; returns K (<0: -lsmc_status). */
int32_t lsmc_parse_basis(const char *expression, int32_t num_factors, int32_t *powers, int32_t capacity);

#ifdef __cplusplus
}
#endif
#endif /* LSMC_B200_H */
