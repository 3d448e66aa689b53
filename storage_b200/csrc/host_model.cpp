// host_model.cpp — see host_model.hpp.  Plain C++ (no CUDA): compiled into liblsmc_b200.so and exercised by the
// CPU test-suite through the exported helper entry points.
#include "host_model.hpp"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstring>
#include <map>
#include <sstream>

namespace lsmc {

namespace {

// Index of the spline segment MathNet's LinearSpline uses for abscissa t: the segment whose left knot is the last
// knot <= t, clamped to [0, n-2] (so values outside the knots extrapolate the end segments).
int left_segment(const double *x, int n, double t) {
  int idx = (int)(std::upper_bound(x, x + n, t) - x) - 1;  // last knot <= t, -1 if none
  if (idx < 0) idx = 0;
  if (idx > n - 2) idx = n - 2;
  return idx;
}

double spline_value(const double *x, const double *y, int n, double t) {
  int k = left_segment(x, n, t);
  double slope = (y[k + 1] - y[k]) / (x[k + 1] - x[k]);
  return y[k] + (t - x[k]) * slope;
}

double horner(const double *c, int len, double z) {
  double acc = c[len - 1];
  for (int i = len - 2; i >= 0; --i) {
    acc *= z;
    acc += c[i];
  }
  return acc;
}

// Root of line through (x1,y1),(x2,y2) at height y (StorageHelper.InterpolateLinearAndSolve, SH.cs:321-330)
double solve_on_segment(double x1, double y1, double x2, double y2, double y) {
  double m = (y2 - y1) / (x2 - x1);
  double c = y1 - m * x1;
  return (y - c) / m;
}

// Safeguarded Newton iteration on [lo, hi]; plays MathNet's RobustNewtonRaphson.TryFindRoot (not vendored).
template <class F, class DF>
bool find_root(F f, DF df, double lo, double hi, double accuracy, int max_iter, int subdivision, double *root) {
  double flo = f(lo), fhi = f(hi);
  if (std::fabs(flo) < accuracy) return *root = lo, true;
  if (std::fabs(fhi) < accuracy) return *root = hi, true;
  if ((flo > 0) == (fhi > 0)) {
    const double width = (hi - lo) / subdivision;
    double a = lo, fa = flo;
    bool bracketed = false;
    for (int i = 1; i <= subdivision && !bracketed; ++i) {
      double b = (i == subdivision) ? hi : lo + i * width, fb = f(b);
      if ((fa > 0) != (fb > 0)) {
        lo = a, hi = b, flo = fa, fhi = fb;
        bracketed = true;
      }
      a = b, fa = fb;
    }
    if (!bracketed) return false;
  }
  double x = 0.5 * (lo + hi), fx = f(x);
  for (int it = 0; it < max_iter; ++it) {
    if (std::fabs(fx) < accuracy) return *root = x, true;
    if ((fx > 0) == (flo > 0))
      lo = x, flo = fx;
    else
      hi = x, fhi = fx;
    double slope = df(x);
    double next = slope != 0.0 ? x - fx / slope : lo - 1.0;
    if (!(next > lo && next < hi)) next = 0.5 * (lo + hi);
    if (std::fabs(next - x) < accuracy * 1e-3) return *root = next, true;
    x = next;
    fx = f(x);
  }
  return false;
}

}  // namespace

// Interpolating polynomial by Householder QR on the Vandermonde matrix (what MathNet's Fit.Polynomial does for
// order = n-1); ascending coefficients.
std::vector<double> fit_polynomial(const double *x, const double *y, int n) {
  std::vector<double> a((size_t)n * n), rhs(y, y + n), v(n);
  for (int i = 0; i < n; ++i) {
    double pw = 1.0;
    for (int j = 0; j < n; ++j, pw *= x[i]) a[(size_t)i * n + j] = pw;
  }
  for (int k = 0; k < n; ++k) {
    double nrm = 0.0;
    for (int i = k; i < n; ++i) nrm += a[(size_t)i * n + k] * a[(size_t)i * n + k];
    nrm = std::sqrt(nrm);
    if (nrm == 0.0) throw Error(LSMC_ERR_ARGUMENT, "Polynomial ratchet fit is singular (repeated inventory pillars?).");
    double alpha = a[(size_t)k * n + k] > 0 ? -nrm : nrm;
    double vv = 0.0;
    for (int i = k; i < n; ++i) {
      v[i] = a[(size_t)i * n + k];
      if (i == k) v[i] -= alpha;
      vv += v[i] * v[i];
    }
    if (vv == 0.0) continue;
    for (int j = k; j < n; ++j) {
      double dot = 0.0;
      for (int i = k; i < n; ++i) dot += v[i] * a[(size_t)i * n + j];
      double f = 2.0 * dot / vv;
      for (int i = k; i < n; ++i) a[(size_t)i * n + j] -= f * v[i];
    }
    double dot = 0.0;
    for (int i = k; i < n; ++i) dot += v[i] * rhs[i];
    double f = 2.0 * dot / vv;
    for (int i = k; i < n; ++i) rhs[i] -= f * v[i];
  }
  std::vector<double> c(n);
  for (int k = n - 1; k >= 0; --k) {
    double s = rhs[k];
    for (int j = k + 1; j < n; ++j) s -= a[(size_t)k * n + j] * c[j];
    c[k] = s / a[(size_t)k * n + k];
  }
  return c;
}

StorageModel::StorageModel(const lsmc_storage &s) : P_(s.num_periods), empty_at_end_(s.must_be_empty_at_end != 0) {
  if (P_ < 0) throw Error(LSMC_ERR_ARGUMENT, "Storage has expired");
  min_inv_.assign(s.min_inventory, s.min_inventory + P_ + 1);
  max_inv_.assign(s.max_inventory, s.max_inventory + P_ + 1);
  if (empty_at_end_) max_inv_[P_] = 0.0;  // CS.cs:435-438
  kind_.assign(s.constraint_kind, s.constraint_kind + P_);
  rbegin_.assign(s.ratchet_begin, s.ratchet_begin + P_);
  rcount_.assign(s.ratchet_count, s.ratchet_count + P_);
  r_inv_.assign(s.ratchet_inventory, s.ratchet_inventory + s.num_ratchet_rows);
  r_min_.assign(s.ratchet_min_rate, s.ratchet_min_rate + s.num_ratchet_rows);
  r_max_.assign(s.ratchet_max_rate, s.ratchet_max_rate + s.num_ratchet_rows);
  auto copy = [&](const double *src, std::vector<double> &dst) { dst.assign(src, src + P_); };
  copy(s.injection_cost, inj_cost_);
  copy(s.injection_cost_df, inj_df_);
  copy(s.withdrawal_cost, wd_cost_);
  copy(s.withdrawal_cost_df, wd_df_);
  copy(s.consumed_on_inject, cons_inj_);
  copy(s.consumed_on_withdraw, cons_wd_);
  copy(s.inventory_loss, loss_);
  copy(s.inventory_cost, inv_cost_);
  copy(s.inventory_cost_df, inv_cost_df_);
  std::memcpy(terminal_, s.terminal_npv_coeff, sizeof(terminal_));
  nr_accuracy_ = s.newton_raphson_accuracy > 0 ? s.newton_raphson_accuracy : 1e-10;
  nr_max_iter_ = s.newton_raphson_max_iterations > 1 ? s.newton_raphson_max_iterations : 100;
  nr_subdivision_ = s.newton_raphson_subdivision > 0 ? s.newton_raphson_subdivision : 20;

  poly_begin_.assign(P_, -1);
  std::map<std::pair<int, int>, int> fitted;  // (begin,count) -> offset, so periods sharing rows share coefficients
  for (int p = 0; p < P_; ++p) {
    if (rcount_[p] < 1 || rbegin_[p] < 0 || rbegin_[p] + rcount_[p] > s.num_ratchet_rows)
      throw Error(LSMC_ERR_ARGUMENT, "Ratchet row range of a period is outside the ratchet table.");
    if (kind_[p] < LSMC_CONSTRAINT_CONSTANT || kind_[p] > LSMC_CONSTRAINT_POLYNOMIAL)
      throw Error(LSMC_ERR_ARGUMENT, "Unknown inject/withdraw constraint kind.");
    if (kind_[p] != LSMC_CONSTRAINT_CONSTANT && rcount_[p] < 2)
      throw Error(LSMC_ERR_ARGUMENT, "Inject/withdraw ranges collection must contain at least two elements.");
    if (kind_[p] != LSMC_CONSTRAINT_POLYNOMIAL) continue;
    auto key = std::make_pair((int)rbegin_[p], (int)rcount_[p]);
    auto it = fitted.find(key);
    if (it == fitted.end()) {
      int off = (int)poly_min_.size();
      const double *xi = &r_inv_[rbegin_[p]];
      std::vector<double> cmin = fit_polynomial(xi, &r_min_[rbegin_[p]], rcount_[p]);
      std::vector<double> cmax = fit_polynomial(xi, &r_max_[rbegin_[p]], rcount_[p]);
      poly_min_.insert(poly_min_.end(), cmin.begin(), cmin.end());
      poly_max_.insert(poly_max_.end(), cmax.begin(), cmax.end());
      it = fitted.emplace(key, off).first;
    }
    poly_begin_[p] = it->second;
  }
}

ConstraintView StorageModel::constraint(int p) const {
  ConstraintView v{};
  v.kind = kind_[p];
  v.count = rcount_[p];
  v.inventory = &r_inv_[rbegin_[p]];
  v.min_rate = &r_min_[rbegin_[p]];
  v.max_rate = &r_max_[rbegin_[p]];
  if (poly_begin_[p] >= 0) {
    v.poly_min = &poly_min_[poly_begin_[p]];
    v.poly_max = &poly_max_[poly_begin_[p]];
    v.poly_len = rcount_[p];
  }
  return v;
}

RateRange StorageModel::constraint_range(int p, double inventory) const {
  ConstraintView c = constraint(p);
  switch (c.kind) {
    case LSMC_CONSTRAINT_CONSTANT:
      return {c.min_rate[0], c.max_rate[0]};
    case LSMC_CONSTRAINT_PIECEWISE_LINEAR:
      return {spline_value(c.inventory, c.min_rate, c.count, inventory), spline_value(c.inventory, c.max_rate, c.count, inventory)};
    case LSMC_CONSTRAINT_STEP: {
      if (inventory < c.inventory[0] || inventory > c.inventory[c.count - 1])
        throw Error(LSMC_ERR_ARGUMENT, "Value of inventory is outside of the ratchet inventory interval.");
      // row whose inventory is the last one <= inventory (Array.BinarySearch semantics, Step…cs:76-78)
      int idx = (int)(std::upper_bound(c.inventory, c.inventory + c.count, inventory) - c.inventory) - 1;
      return {c.min_rate[idx], c.max_rate[idx]};
    }
    default:
      return {horner(c.poly_min, c.poly_len, inventory), horner(c.poly_max, c.poly_len, inventory)};
  }
}

RateRange StorageModel::inject_withdraw_range(int p, double inventory) const {
  double lo = min_inventory(p);
  if (inventory < lo) {
    std::ostringstream m;
    m << "Inventory of " << inventory << " is below minimum allowed value of " << lo << " during period " << p << ".";
    throw Error(LSMC_ERR_ARGUMENT, m.str());
  }
  double hi = max_inventory(p);
  if (inventory > hi) {
    std::ostringstream m;
    m << "Inventory of " << inventory << " above maximum allowed value of " << hi << " during period " << p << ".";
    throw Error(LSMC_ERR_ARGUMENT, m.str());
  }
  if (p >= P_) return {0.0, 0.0};
  return constraint_range(p, inventory);
}

double StorageModel::space_upper_bound(int p, double next_lo, double next_hi) const {
  const ConstraintView c = constraint(p);
  const double cur_min = min_inventory(p), cur_max = max_inventory(p), keep = 1 - loss_[p];
  if (c.kind == LSMC_CONSTRAINT_CONSTANT) return std::min((next_hi - c.min_rate[0]) / keep, cur_max);

  const bool poly = c.kind == LSMC_CONSTRAINT_POLYNOMIAL;
  RateRange at_max = constraint_range(p, cur_max);
  // reachable from the current maximum?  (the polynomial class omits the loss factor here, Polynomial…cs:94-95)
  double base = poly ? cur_max : cur_max * keep;
  double reach_hi = base + at_max.hi, reach_lo = base + at_max.lo;
  if (reach_lo <= next_hi && next_lo <= reach_hi) return cur_max;

  if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
    double up_inv = c.inventory[c.count - 1], up_after = reach_lo;
    for (int i = c.count - 2; i >= 0; --i) {
      double lo_inv = c.inventory[i], lo_after = lo_inv * keep + c.min_rate[i];
      if (lo_after <= next_hi && next_hi <= up_after) return solve_on_segment(lo_inv, lo_after, up_inv, up_after, next_hi);
      up_after = lo_after;
      up_inv = lo_inv;
    }
    throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
  }
  if (c.kind == LSMC_CONSTRAINT_STEP) {
    bool any = false;
    double best = 0.0;
    for (int i = 0; i + 1 < c.count; ++i) {  // keeps the highest solution (Step…cs:113-116)
      double rate = c.min_rate[i];
      double lo_after = c.inventory[i] * keep + rate, up_after = c.inventory[i + 1] * keep + rate;
      if (lo_after <= next_hi && next_hi <= up_after) {
        best = solve_on_segment(c.inventory[i], lo_after, c.inventory[i + 1], up_after, next_hi);
        any = true;
      }
    }
    if (!any) throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
    return best;
  }
  // polynomial: solve inv*(1-loss) + minRate(inv) = next_hi on [cur_min, cur_max] (Polynomial…cs:104-119)
  std::vector<double> deriv(std::max(c.poly_len - 1, 1), 0.0);
  for (int i = 1; i < c.poly_len; ++i) deriv[i - 1] = c.poly_min[i] * i;
  auto f = [&](double x) { return x * keep + horner(c.poly_min, c.poly_len, x) - next_hi; };
  auto df = [&](double x) { return keep + horner(deriv.data(), (int)deriv.size(), x); };
  double root;
  if (!find_root(f, df, cur_min, cur_max, nr_accuracy_, nr_max_iter_, nr_subdivision_, &root))
    throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS,
                "Cannot solve for the current period inventory space upper bound. Try changing Newton Raphson parameters.");
  if (root < cur_min) throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Inventory constraints cannot be satisfied.");
  return std::min(root, cur_max);
}

double StorageModel::space_lower_bound(int p, double next_lo, double next_hi) const {
  const ConstraintView c = constraint(p);
  const double cur_min = min_inventory(p), cur_max = max_inventory(p), keep = 1 - loss_[p];
  if (c.kind == LSMC_CONSTRAINT_CONSTANT) return std::max((next_lo - c.max_rate[0]) / keep, cur_min);

  const bool poly = c.kind == LSMC_CONSTRAINT_POLYNOMIAL;
  RateRange at_min = constraint_range(p, cur_min);
  double base = poly ? cur_min : cur_min * keep;
  double reach_hi = base + at_min.hi, reach_lo = base + at_min.lo;
  if (reach_lo <= next_hi && next_lo <= reach_hi) return cur_min;

  if (c.kind == LSMC_CONSTRAINT_PIECEWISE_LINEAR) {
    double lo_inv = c.inventory[0], lo_after = reach_hi;
    for (int i = 1; i < c.count; ++i) {
      double up_inv = c.inventory[i], up_after = up_inv * keep + c.max_rate[i];
      if (lo_after <= next_lo && next_lo <= up_after) return solve_on_segment(lo_inv, lo_after, up_inv, up_after, next_lo);
      lo_after = up_after;
      lo_inv = up_inv;
    }
    throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
  }
  if (c.kind == LSMC_CONSTRAINT_STEP) {
    bool any = false;
    double best = 0.0;
    for (int i = c.count - 2; i >= 0; --i) {  // keeps the lowest solution (Step…cs:156-159)
      double rate = c.max_rate[i];
      double lo_after = c.inventory[i] * keep + rate, up_after = c.inventory[i + 1] * keep + rate;
      if (lo_after <= next_lo && next_lo <= up_after) {
        best = solve_on_segment(c.inventory[i], lo_after, c.inventory[i + 1], up_after, next_lo);
        any = true;
      }
    }
    if (!any) throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
    return best;
  }
  std::vector<double> deriv(std::max(c.poly_len - 1, 1), 0.0);
  for (int i = 1; i < c.poly_len; ++i) deriv[i - 1] = c.poly_max[i] * i;
  auto f = [&](double x) { return x * keep + horner(c.poly_max, c.poly_len, x) - next_lo; };
  auto df = [&](double x) { return keep + horner(deriv.data(), (int)deriv.size(), x); };
  double root;
  if (!find_root(f, df, cur_min, cur_max, nr_accuracy_, nr_max_iter_, nr_subdivision_, &root))
    throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS,
                "Cannot solve for the current period inventory space lower bound. Try changing Newton Raphson parameters.");
  if (root > cur_max) throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Inventory constraints cannot be satisfied.");
  return std::max(root, cur_min);
}

double StorageModel::decision_cost_npv(int p, double volume) const {
  if (volume > 0.0) return (inj_cost_[p] * volume) * inj_df_[p];
  return (wd_cost_[p] * std::fabs(-volume)) * wd_df_[p];
}

double StorageModel::decision_consumed(int p, double volume) const {
  return (volume > 0.0 ? cons_inj_[p] : cons_wd_[p]) * std::fabs(volume);
}

double StorageModel::inventory_cost_npv(int p, double inventory) const { return (inventory * inv_cost_[p]) * inv_cost_df_[p]; }

double StorageModel::terminal_npv(double price, double inventory) const {
  return ((terminal_[3] * (price * inventory) + terminal_[1] * price) + terminal_[2] * inventory) + terminal_[0];
}

void calculate_inventory_space(const StorageModel &st, double starting_inventory, std::vector<double> &lo, std::vector<double> &hi) {
  const int P = st.num_periods();
  lo.assign(P, 0.0);
  hi.assign(P, 0.0);
  if (P == 0) return;
  // forward sweep: what can be reached from the starting inventory (SH.cs:54-74)
  double reach_lo = starting_inventory, reach_hi = starting_inventory;
  for (int p = 0; p < P; ++p) {
    const double loss = st.loss_fraction(p);
    double down = st.inject_withdraw_range(p, reach_lo).lo;
    reach_lo = std::max(reach_lo - loss * reach_lo + down, st.min_inventory(p + 1));
    double up = st.inject_withdraw_range(p, reach_hi).hi;
    reach_hi = std::min(reach_hi - loss * reach_hi + up, st.max_inventory(p + 1));
    lo[p] = reach_lo;
    hi[p] = reach_hi;
  }
  // backward sweep: from where the terminal condition can still be met (SH.cs:76-91), intersected (SH.cs:95-104)
  double need_lo = st.must_be_empty_at_end() ? 0.0 : st.min_inventory(P);
  double need_hi = st.must_be_empty_at_end() ? 0.0 : st.max_inventory(P);
  for (int i = P - 1; i >= 0; --i) {
    if (i < P - 1) {
      double up = st.space_upper_bound(i + 1, need_lo, need_hi);
      double down = st.space_lower_bound(i + 1, need_lo, need_hi);
      need_hi = up;
      need_lo = down;
    }
    hi[i] = std::min(hi[i], need_hi);
    lo[i] = std::max(lo[i], need_lo);
    if (lo[i] > hi[i]) throw Error(LSMC_ERR_INVENTORY_CONSTRAINTS, "Inventory constraints cannot be fulfilled.");
  }
}

std::vector<double> grid_points(double lower, double upper, double spacing) {
  if (lower > upper) throw Error(LSMC_ERR_ARGUMENT, "Parameter stateSpaceLowerBound value cannot be above parameter stateSpaceUpperBound value");
  if (!(spacing > 0.0)) throw Error(LSMC_ERR_ARGUMENT, "Grid spacing must be positive");
  std::vector<double> pts{lower};
  for (double x = lower; x < upper;) {  // repeated addition, like the reference's iterator
    x += spacing;
    pts.push_back(x < upper ? x : upper);
  }
  return pts;
}

int bang_bang_decision_set(RateRange range, double inventory, double inventory_loss, double next_min, double next_max,
                           double tolerance, int extra, double *out) {
  if (next_min > next_max)
    throw Error(LSMC_ERR_ARGUMENT, "Parameter nextStepMinInventory value cannot be higher than parameter nextStepMaxInventory value.");
  if (extra < 0) throw Error(LSMC_ERR_ARGUMENT, "Parameter numExtraDecisions must be non-negative.");
  if (2 * extra + 3 > LSMC_MAX_DECISIONS) throw Error(LSMC_ERR_ARGUMENT, "Too many extra decisions for LSMC_MAX_DECISIONS.");
  const char *infeasible =
      "Inventory constraints cannot be fulfilled. This could potentially be fixed by increasing the numerical tolerance.";
  const double base = inventory - inventory_loss;

  double withdraw;
  const double lowest = range.lo + base;  // inventory after the largest withdrawal
  if (lowest > next_max) {
    if (!(lowest - next_max < tolerance)) throw Error(LSMC_ERR_ARGUMENT, infeasible);
    withdraw = next_max - base;
  } else {
    withdraw = lowest > next_min ? range.lo : next_min - base;
  }
  double inject;
  const double highest = range.hi + base;  // inventory after the largest injection
  if (highest < next_min) {
    if (!(next_min - highest < tolerance)) throw Error(LSMC_ERR_ARGUMENT, infeasible);
    inject = next_min - base;
  } else {
    inject = highest < next_max ? range.hi : next_max - base;
  }

  auto fill_between = [extra](double from, double to, double *dst) {
    const double step = (to - from) / (extra + 1);
    for (int i = 0; i < extra; ++i) dst[i] = (i + 1) * step + from;
  };
  int n = 0;
  out[n++] = withdraw;
  if (withdraw >= 0.0 || inject <= 0.0) {  // no "do nothing" decision
    fill_between(withdraw, inject, out + n);
    n += extra;
  } else {
    fill_between(withdraw, 0, out + n);
    n += extra;
    out[n++] = 0.0;
    fill_between(0, inject, out + n);
    n += extra;
  }
  out[n++] = inject;
  return n;
}

void bisect_inventory_space(const double *grid, int n, double inventory, double tol, int *lower, int *upper) {
  auto close = [tol](double a, double b) { return std::fabs(a - b) <= tol; };
  if (n == 1 && close(inventory, grid[0])) {
    *lower = *upper = 0;
    return;
  }
  int left = 0, right = n - 1;
  const int top = right;
  while (right > left) {
    const int mid = (left + right) / 2;
    if (close(inventory, grid[mid])) {
      *lower = *upper = mid;
      return;
    }
    if (grid[mid] > inventory) {
      right = mid;
      continue;
    }
    if (inventory <= grid[mid + 1]) {
      *lower = mid;
      *upper = mid + 1;
      return;
    }
    if (close(inventory, grid[mid + 1])) {
      *lower = *upper = mid + 1;
      return;
    }
    if (mid + 1 == top) break;
    left = mid;
  }
  throw Error(LSMC_ERR_ARGUMENT, "Inventory is outside of inventoryGrid bounds.");
}

// ---- basis expression parser ----------------------------------------------------------------------------------
// Grammar accepted by the reference (BasisFunctionsBuilder.cs:90-131): monomials separated by '+'; a monomial is "1"
// or a '*'-product of terms; a term is s | S | x<i>, optionally followed by **<power>.
int parse_basis(const char *expression, int num_factors, int32_t *powers, int capacity) {
  if (!expression) throw Error(LSMC_ERR_ARGUMENT, "basisFunctionExpression is null");
  const int width = 1 + num_factors;
  std::vector<std::string> monomials;
  {
    std::string cur;
    for (const char *c = expression;; ++c) {
      if (*c == '+' || *c == '\0') {
        size_t b = cur.find_first_not_of(" \t"), e = cur.find_last_not_of(" \t");
        monomials.push_back(b == std::string::npos ? std::string() : cur.substr(b, e - b + 1));
        cur.clear();
        if (*c == '\0') break;
      } else {
        cur.push_back(*c);
      }
    }
  }
  for (size_t i = 0; i < monomials.size(); ++i)
    for (size_t j = i + 1; j < monomials.size(); ++j)
      if (monomials[i] == monomials[j]) throw Error(LSMC_ERR_ARGUMENT, "Basis function expression contains repeated monomials.");
  if ((int)monomials.size() > capacity) throw Error(LSMC_ERR_ARGUMENT, "Too many basis functions.");
  int k = 0;
  for (const std::string &m : monomials) {
    int32_t *row = powers + (size_t)k * width;
    std::fill(row, row + width, 0);
    if (m.empty()) throw Error(LSMC_ERR_ARGUMENT, "Basis function expression contains an empty monomial.");
    if (m != "1") {
      size_t pos = 0;
      auto skip_ws = [&] { while (pos < m.size() && std::isspace((unsigned char)m[pos])) ++pos; };
      auto read_uint = [&]() -> int {
        skip_ws();
        if (pos >= m.size() || !std::isdigit((unsigned char)m[pos])) throw Error(LSMC_ERR_ARGUMENT, "Cannot parse monomial '" + m + "'.");
        int v = 0;
        while (pos < m.size() && std::isdigit((unsigned char)m[pos])) v = v * 10 + (m[pos++] - '0');
        return v;
      };
      for (;;) {
        skip_ws();
        if (pos >= m.size()) throw Error(LSMC_ERR_ARGUMENT, "Cannot parse monomial '" + m + "'.");
        int column;
        if (m[pos] == 's' || m[pos] == 'S') {
          column = 0;
          ++pos;
        } else if (m[pos] == 'x' || m[pos] == 'X') {
          ++pos;
          int f = read_uint();
          if (f >= num_factors) throw Error(LSMC_ERR_ARGUMENT, "Monomial '" + m + "' uses a factor the simulation does not have.");
          column = 1 + f;
        } else {
          throw Error(LSMC_ERR_ARGUMENT, "Cannot parse monomial '" + m + "'.");
        }
        int power = 1;
        skip_ws();
        if (pos + 1 < m.size() && m[pos] == '*' && m[pos + 1] == '*') {
          pos += 2;
          power = read_uint();
        }
        row[column] += power;
        skip_ws();
        if (pos >= m.size()) break;
        if (m[pos] != '*') throw Error(LSMC_ERR_ARGUMENT, "Cannot parse monomial '" + m + "'.");
        ++pos;
      }
    }
    ++k;
  }
  return k;
}

// ---- backward tables -----------------------------------------------------------------------------------------
ValuationTables build_tables(const StorageModel &st, const lsmc_problem &pb) {
  ValuationTables t;
  const int P = st.num_periods();
  t.P = P;
  std::vector<double> lo, hi;
  calculate_inventory_space(st, pb.inventory, lo, hi);
  t.space_lo.assign(P + 1, 0.0);
  t.space_hi.assign(P + 1, 0.0);
  for (int i = 0; i < P; ++i) {
    t.space_lo[i + 1] = lo[i];
    t.space_hi[i + 1] = hi[i];
  }
  t.periods.resize(P + 1);
  t.periods[0].grid = {pb.inventory};  // LSV.cs:216-217
  for (int p = 1; p <= P; ++p) t.periods[p].grid = grid_points(t.space_lo[p], t.space_hi[p], pb.grid_spacing);
  t.max_decisions = pb.extra_decisions > 0 ? 2 * pb.extra_decisions + 3 : 3;

  for (int p = 0; p < P; ++p) {
    PeriodTables &pt = t.periods[p];
    const std::vector<double> &next = t.periods[p + 1].grid;
    const int G = (int)pt.grid.size(), Gn = (int)next.size();
    t.max_grid = std::max(t.max_grid, std::max(G, Gn));
    t.grid_point_steps += G;
    pt.inventory_cost.resize(G);
    pt.num_decisions.resize(G);
    pt.decisions.assign((size_t)G * t.max_decisions, DecisionEntry{0, 0, 0, 1.0, 0.0, 0, 0});
    double volumes[LSMC_MAX_DECISIONS];
    for (int g = 0; g < G; ++g) {
      const double inventory = pt.grid[g];
      RateRange range = st.inject_withdraw_range(p, inventory);
      const double loss = st.loss_fraction(p) * inventory;
      const int D = bang_bang_decision_set(range, inventory, loss, t.space_lo[p + 1], t.space_hi[p + 1], pb.numerical_tolerance,
                                           pb.extra_decisions, volumes);
      pt.num_decisions[g] = D;
      pt.inventory_cost[g] = st.inventory_cost_npv(p, inventory);
      for (int d = 0; d < D; ++d) {
        DecisionEntry &e = pt.decisions[(size_t)g * t.max_decisions + d];
        e.volume = volumes[d];
        e.cost_npv = st.decision_cost_npv(p, volumes[d]);
        e.consumed = st.decision_consumed(p, volumes[d]);
        const double target = inventory + volumes[d] - loss;
        // The reference scans the next grid upwards and stops at the first point that is within 1e-8 of the target or
        // above it (LSV.cs:267-305).  Equivalent: the first point > target - 1e-8.
        int j = (int)(std::lower_bound(next.begin(), next.end(), target,
                                       [](double point, double tgt) { return !(std::fabs(point - tgt) < 1E-8) && !(point > tgt); }) -
                      next.begin());
        if (j == Gn) throw Error(LSMC_ERR_ARGUMENT, "Inventory after a decision is above the next period's inventory grid.");
        if (std::fabs(next[j] - target) < 1E-8) {
          e.lower = e.upper = j;
          e.lower_w = 1.0;
          e.upper_w = 0.0;
        } else {
          if (j == 0) throw Error(LSMC_ERR_ARGUMENT, "Inventory after a decision is below the next period's inventory grid.");
          e.lower = j - 1;
          e.upper = j;
          e.lower_w = (next[j] - target) / (next[j] - next[j - 1]);
          e.upper_w = 1.0 - e.lower_w;
        }
      }
    }
  }
  t.max_grid = std::max(t.max_grid, (int)t.periods[P].grid.size());
  return t;
}

}  // namespace lsmc
