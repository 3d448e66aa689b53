// host_model.hpp — sim-independent host logic of the product library: the storage entity over the flat descriptor,
// inventory space, grids, bang-bang decision sets, grid bisection, and the per-(period, grid point, decision) tables
// the backward kernel consumes.  Everything here is O(T*G) work done once per valuation in lsmc_create.
//
// Reference semantics (citations relative to /root/reference/src/Cmdty.Storage/):
//   StorageHelper.cs:39-107,109-204,280-316; StorageEntity/CmdtyStorage.cs:86-169;
//   InjectWithdrawConstraints/*.cs; StateSpaceGridCalc/FixedSpacingStateSpaceGridCalc.cs:45-63;
//   LsmcValuation/LsmcStorageValuation.cs:215-306 (what is tabulated per grid point).
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/lsmc_b200.h"

namespace lsmc {

struct Error : std::runtime_error {
  lsmc_status status;
  Error(lsmc_status s, const std::string &msg) : std::runtime_error(msg), status(s) {}
};

struct RateRange {
  double lo;  // MinInjectWithdrawRate
  double hi;  // MaxInjectWithdrawRate
};

// One period's inject/withdraw constraint in the form both host and device evaluate.
struct ConstraintView {
  int kind;
  int count;
  const double *inventory;
  const double *min_rate;
  const double *max_rate;
  // POLYNOMIAL only: ascending coefficients of the fitted min/max rate polynomials
  const double *poly_min;
  const double *poly_max;
  int poly_len;
};

// Deep copy of lsmc_storage with the derived data (polynomial coefficients, MustBeEmptyAtEnd override) resolved.
class StorageModel {
 public:
  explicit StorageModel(const lsmc_storage &s);

  int num_periods() const { return P_; }
  bool must_be_empty_at_end() const { return empty_at_end_; }
  double min_inventory(int p) const { return min_inv_[p > P_ ? P_ : p]; }
  double max_inventory(int p) const { return max_inv_[p > P_ ? P_ : p]; }  // override already applied at P
  double loss_fraction(int p) const { return loss_[p]; }
  ConstraintView constraint(int p) const;

  RateRange constraint_range(int p, double inventory) const;        // IInjectWithdrawConstraint.GetInjectWithdrawRange
  RateRange inject_withdraw_range(int p, double inventory) const;   // CmdtyStorage.GetInjectWithdrawRange (CS.cs:86-100)
  double space_upper_bound(int p, double next_lo, double next_hi) const;  // CS.cs:132-141
  double space_lower_bound(int p, double next_lo, double next_hi) const;  // CS.cs:143-152
  double decision_cost_npv(int p, double volume) const;             // LSV.cs:765-774
  double decision_consumed(int p, double volume) const;             // LSV.cs:757-763
  double inventory_cost_npv(int p, double inventory) const;         // LSV.cs:247-248
  double terminal_npv(double price, double inventory) const;        // CS.cs:154-157

  // raw per-period parameter arrays, uploaded to the device for the forward pass
  std::vector<double> min_inv_, max_inv_;
  std::vector<int32_t> kind_, rbegin_, rcount_;
  std::vector<double> r_inv_, r_min_, r_max_;
  std::vector<int32_t> poly_begin_;            // per period: offset into poly_min_/poly_max_ (-1 if not polynomial)
  std::vector<double> poly_min_, poly_max_;    // ascending coefficients, rcount_[p] per polynomial constraint
  std::vector<double> inj_cost_, inj_df_, wd_cost_, wd_df_, cons_inj_, cons_wd_, loss_, inv_cost_, inv_cost_df_;
  double terminal_[4];
  double nr_accuracy_;
  int nr_max_iter_, nr_subdivision_;

 private:
  int P_;
  bool empty_at_end_;
};

// StorageHelper.CalculateInventorySpace: ranges for periods 1..P (index i = period i+1).
void calculate_inventory_space(const StorageModel &st, double starting_inventory, std::vector<double> &lo, std::vector<double> &hi);
// FixedSpacingStateSpaceGridCalc.GetGridPoints
std::vector<double> grid_points(double lower, double upper, double spacing);
// StorageHelper.CalculateBangBangDecisionSet: returns the number of decisions written to out[LSMC_MAX_DECISIONS]
int bang_bang_decision_set(RateRange range, double inventory, double inventory_loss, double next_min, double next_max,
                           double tolerance, int extra_decisions, double *out);
// StorageHelper.BisectInventorySpace
void bisect_inventory_space(const double *grid, int n, double inventory, double tolerance, int *lower, int *upper);
// Roslyn-free BasisFunctionsBuilder.Parse; returns the number of monomials
int parse_basis(const char *expression, int num_factors, int32_t *powers, int capacity);
// interpolating polynomial through (x_i, y_i): ascending coefficients (MathNet Fit.Polynomial with order n-1)
std::vector<double> fit_polynomial(const double *x, const double *y, int n);

// ---- backward tables -----------------------------------------------------------------------------------------
// One decision of one grid point of one period (LSV.cs:255-306): everything that does not depend on the path.
struct DecisionEntry {
  double volume;      // decisionVolume
  double cost_npv;    // injectWithdrawCostNpvs[d]
  double consumed;    // cmdtyUsedForInjectWithdrawVolume[d]
  double lower_w;     // lowerWeight (1 when the target hits a grid point)
  double upper_w;     // upperWeight (0 when the target hits a grid point)
  int32_t lower;      // index into the next period's grid
  int32_t upper;      // == lower when the target hits a grid point within 1e-8 (LSV.cs:270)
};

struct PeriodTables {
  std::vector<double> grid;             // inventory grid of this period (LSV.cs:215-223)
  std::vector<double> inventory_cost;   // [G] inventoryCostNpv
  std::vector<int32_t> num_decisions;   // [G]
  std::vector<DecisionEntry> decisions; // [G][max_decisions]
};

struct ValuationTables {
  int P = 0;
  int max_decisions = 0;
  int max_grid = 0;
  std::vector<double> space_lo, space_hi;  // [P+1], index = period (entry 0 unused)
  std::vector<PeriodTables> periods;       // [P+1]; entry P holds only the grid
  int64_t grid_point_steps = 0;            // sum over decision periods of G_p (the metric's "inventory points")
};

ValuationTables build_tables(const StorageModel &st, const lsmc_problem &pb);

}  // namespace lsmc
