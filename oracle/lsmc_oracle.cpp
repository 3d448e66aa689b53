// lsmc_oracle.cpp — CPU restatement of cmdty/storage's LSMC valuation path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the parity oracle: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.  Nothing under storage_b200/ links, imports or calls it.
//
// The reference (C#, .NET) cannot be compiled in this image (no dotnet/mono/csc), so this is a port, written
// operation-for-operation from the reference sources.  Citations are relative to
// /root/reference/src/Cmdty.Storage/ :  LSV.cs = LsmcValuation/LsmcStorageValuation.cs, SH.cs = StorageHelper.cs,
// CS.cs = StorageEntity/CmdtyStorage.cs.
//
// Parity pinning: the sim-independent helpers are pinned by the reference's own exact-equality unit tests
// (tests/test_oracle_kat.py ports them).  The regression arithmetic lives in the un-vendored MathNet.Numerics
// (QR / Inverse / Multiply, LSV.cs:193-205) for which the reference holds no isolated test vector, and the
// end-to-end goldens need the .NET Mersenne-Twister normal stream of the un-vendored Cmdty.Core 0.2.0:
// for Calculate() as a whole parity is therefore UNPINNED against reference outputs.
//
// Two regression modes:
//   0  "qr"    : thin Householder QR -> R^-1 -> pinv = R^-1 Q^T -> beta = pinv*y -> X*beta, all in double, i.e. the
//                reference's algorithm (LSV.cs:193-206) with a fixed (sequential) summation order.
//   1  "exact" : the same least-squares fit evaluated in double-double arithmetic and rounded once, i.e. the value
//                every correct double implementation of LSV.cs:193-206 approximates.  The GPU is checked against
//                this mode: argmax decisions are only well defined against correctly rounded continuation values.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/lsmc_b200.h"

namespace {

struct OracleError : std::runtime_error {
  int code;
  OracleError(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

thread_local std::string g_last_error;

// ------------------------------------------------------------------------------------------------
// double-double arithmetic (error-free transformations; used by the "exact" regression mode only)
// ------------------------------------------------------------------------------------------------
struct dd {
  double hi, lo;
};
inline dd two_sum(double a, double b) {
  double s = a + b, bb = s - a;
  return {s, (a - (s - bb)) + (b - bb)};
}
inline dd quick_two_sum(double a, double b) {
  double s = a + b;
  return {s, b - (s - a)};
}
inline dd two_prod(double a, double b) {
  double p = a * b;
  return {p, std::fma(a, b, -p)};
}
inline dd dd_add(dd a, dd b) {
  dd s = two_sum(a.hi, b.hi), t = two_sum(a.lo, b.lo);
  s.lo += t.hi;
  s = quick_two_sum(s.hi, s.lo);
  s.lo += t.lo;
  return quick_two_sum(s.hi, s.lo);
}
inline dd dd_neg(dd a) { return {-a.hi, -a.lo}; }
inline dd dd_sub(dd a, dd b) { return dd_add(a, dd_neg(b)); }
inline dd dd_mul(dd a, dd b) {
  dd p = two_prod(a.hi, b.hi);
  p.lo += a.hi * b.lo + a.lo * b.hi;
  return quick_two_sum(p.hi, p.lo);
}
inline dd dd_mul_d(dd a, double b) {
  dd p = two_prod(a.hi, b);
  p.lo += a.lo * b;
  return quick_two_sum(p.hi, p.lo);
}
inline dd dd_div(dd a, dd b) {
  double q1 = a.hi / b.hi;
  dd r = dd_sub(a, dd_mul_d(b, q1));
  double q2 = r.hi / b.hi;
  r = dd_sub(r, dd_mul_d(b, q2));
  double q3 = r.hi / b.hi;
  dd q = quick_two_sum(q1, q2);
  return dd_add(q, {q3, 0.0});
}
inline dd dd_sqrt(dd a) {
  if (a.hi <= 0.0) return {0.0, 0.0};
  double x = 1.0 / std::sqrt(a.hi);
  double ax = a.hi * x;
  dd ax2 = two_prod(ax, ax);
  dd diff = dd_sub(a, ax2);
  return dd_add({ax, 0.0}, {diff.hi * (x * 0.5), 0.0});
}

// ------------------------------------------------------------------------------------------------
// Storage entity: ICmdtyStorage<T> over the structured descriptor (CS.cs:86-169 + builder lambdas :203-568)
// ------------------------------------------------------------------------------------------------
struct Range {
  double min_rate, max_rate;
};

// MathNet LinearSpline.InterpolateSorted + Interpolate (used by PiecewiseLinear…cs:62-70): c0=y_k, c1=slope_k,
// segment index = clamp(BinarySearch-left, 0, n-2), value = y_k + (t - x_k)*slope_k (extrapolates linearly).
double linear_spline(const double *x, const double *y, int n, double t) {
  int lo = 0, hi = n - 1, idx = -1;
  // Array.BinarySearch semantics: exact match index, else ~(index of first larger element)
  while (lo <= hi) {
    int mid = lo + ((hi - lo) >> 1);
    if (x[mid] == t) {
      idx = mid;
      break;
    }
    if (x[mid] < t)
      lo = mid + 1;
    else
      hi = mid - 1;
  }
  if (idx < 0) idx = lo - 1;  // ~(~lo) - 1
  idx = std::min(std::max(idx, 0), n - 2);
  double slope = (y[idx + 1] - y[idx]) / (x[idx + 1] - x[idx]);
  return y[idx] + (t - x[idx]) * slope;
}

double interpolate_linear_and_solve(double x1, double y1, double x2, double y2, double y) {  // SH.cs:321-330
  double gradient = (y2 - y1) / (x2 - x1);
  double constant = y1 - gradient * x1;
  return (y - constant) / gradient;
}

// Polynomial through the ratchet points (Polynomial…cs:62-73).  MathNet's Fit.Polynomial solves the Vandermonde
// least-squares system by QR; with order = points-1 the fit is the interpolating polynomial.  Solved here by
// Householder QR on the Vandermonde matrix in double (MathNet is not vendored: coefficients agree to rounding only).
std::vector<double> poly_fit(const double *x, const double *y, int n) {
  int m = n;
  std::vector<double> a(n * m), b(y, y + n);
  for (int i = 0; i < n; i++) {
    double p = 1.0;
    for (int j = 0; j < m; j++) {
      a[i * m + j] = p;
      p *= x[i];
    }
  }
  for (int k = 0; k < m; k++) {
    double norm = 0.0;
    for (int i = k; i < n; i++) norm += a[i * m + k] * a[i * m + k];
    norm = std::sqrt(norm);
    if (norm == 0.0) throw OracleError(LSMC_ERR_ARGUMENT, "Polynomial fit is singular.");
    double alpha = a[k * m + k] > 0 ? -norm : norm;
    std::vector<double> v(n, 0.0);
    for (int i = k; i < n; i++) v[i] = a[i * m + k];
    v[k] -= alpha;
    double vnorm2 = 0.0;
    for (int i = k; i < n; i++) vnorm2 += v[i] * v[i];
    if (vnorm2 > 0.0) {
      for (int j = k; j < m; j++) {
        double dot = 0.0;
        for (int i = k; i < n; i++) dot += v[i] * a[i * m + j];
        double f = 2.0 * dot / vnorm2;
        for (int i = k; i < n; i++) a[i * m + j] -= f * v[i];
      }
      double dot = 0.0;
      for (int i = k; i < n; i++) dot += v[i] * b[i];
      double f = 2.0 * dot / vnorm2;
      for (int i = k; i < n; i++) b[i] -= f * v[i];
    }
  }
  std::vector<double> c(m);
  for (int k = m - 1; k >= 0; k--) {
    double s = b[k];
    for (int j = k + 1; j < m; j++) s -= a[k * m + j] * c[j];
    c[k] = s / a[k * m + k];
  }
  return c;
}
double poly_eval(const std::vector<double> &c, double z) {  // MathNet Polynomial.Evaluate: Horner
  double sum = c.back();
  for (int i = (int)c.size() - 2; i >= 0; i--) {
    sum *= z;
    sum += c[i];
  }
  return sum;
}
std::vector<double> poly_diff(const std::vector<double> &c) {
  std::vector<double> d(std::max<size_t>(c.size() - 1, 1), 0.0);
  for (size_t i = 1; i < c.size(); i++) d[i - 1] = c[i] * (double)i;
  return d;
}

// Stand-in for MathNet RobustNewtonRaphson.TryFindRoot (Polynomial…cs:109,144; MathNet is not vendored): Newton steps
// safeguarded by bisection on a sign-change bracket found by scanning `subdivision` sub-intervals.
bool robust_newton(const std::function<double(double)> &f, const std::function<double(double)> &df, double lo, double hi,
                   double accuracy, int max_iter, int subdivision, double *root) {
  double flo = f(lo), fhi = f(hi);
  if (std::fabs(flo) < accuracy) {
    *root = lo;
    return true;
  }
  if (std::fabs(fhi) < accuracy) {
    *root = hi;
    return true;
  }
  if ((flo > 0) == (fhi > 0)) {  // no sign change at the ends: scan sub-intervals for one
    bool found = false;
    double step = (hi - lo) / subdivision, a = lo, fa = flo;
    for (int i = 1; i <= subdivision; i++) {
      double b = (i == subdivision) ? hi : lo + i * step, fb = f(b);
      if ((fa > 0) != (fb > 0)) {
        lo = a;
        hi = b;
        flo = fa;
        fhi = fb;
        found = true;
        break;
      }
      a = b;
      fa = fb;
    }
    if (!found) return false;
  }
  double x = 0.5 * (lo + hi), fx = f(x);
  for (int it = 0; it < max_iter; it++) {
    if (std::fabs(fx) < accuracy) {
      *root = x;
      return true;
    }
    if ((fx > 0) == (flo > 0)) {
      lo = x;
      flo = fx;
    } else {
      hi = x;
      fhi = fx;
    }
    double d = df(x), xn = (d != 0.0) ? x - fx / d : lo - 1.0;
    if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
    if (std::fabs(xn - x) < accuracy * 1e-3) {
      *root = xn;
      return true;
    }
    x = xn;
    fx = f(x);
  }
  return false;
}

struct Storage {
  const lsmc_storage &s;
  int P;
  explicit Storage(const lsmc_storage &st) : s(st), P(st.num_periods) {}

  double min_inventory(int p) const { return s.min_inventory[std::min(p, P)]; }
  double max_inventory(int p) const {  // CS.cs:435-438: MustBeEmptyAtEnd forces 0 from the end period on
    if (s.must_be_empty_at_end && p >= P) return 0.0;
    return s.max_inventory[std::min(p, P)];
  }
  double loss_pct(int p) const { return s.inventory_loss[p]; }

  const double *rinv(int p) const { return s.ratchet_inventory + s.ratchet_begin[p]; }
  const double *rmin(int p) const { return s.ratchet_min_rate + s.ratchet_begin[p]; }
  const double *rmax(int p) const { return s.ratchet_max_rate + s.ratchet_begin[p]; }
  int rcount(int p) const { return s.ratchet_count[p]; }

  // IInjectWithdrawConstraint.GetInjectWithdrawRange for the four implementations
  Range constraint_range(int p, double inventory) const {
    const double *xi = rinv(p), *mn = rmin(p), *mx = rmax(p);
    int n = rcount(p);
    switch (s.constraint_kind[p]) {
      case LSMC_CONSTRAINT_CONSTANT:  // Constant…cs:45-48
        return {mn[0], mx[0]};
      case LSMC_CONSTRAINT_PIECEWISE_LINEAR:  // PiecewiseLinear…cs:67-72 (max first, then min)
        return {linear_spline(xi, mn, n, inventory), linear_spline(xi, mx, n, inventory)};
      case LSMC_CONSTRAINT_STEP: {  // Step…cs:72-79
        if (inventory < xi[0] || inventory > xi[n - 1])
          throw OracleError(LSMC_ERR_ARGUMENT, "Value of inventory is outside of the ratchet interval.");
        int lo = 0, hi = n - 1, idx = -1;
        while (lo <= hi) {
          int mid = lo + ((hi - lo) >> 1);
          if (xi[mid] == inventory) {
            idx = mid;
            break;
          }
          if (xi[mid] < inventory)
            lo = mid + 1;
          else
            hi = mid - 1;
        }
        if (idx < 0) idx = lo - 1;
        return {mn[idx], mx[idx]};
      }
      case LSMC_CONSTRAINT_POLYNOMIAL: {  // Polynomial…cs:80-85
        std::vector<double> cmax = poly_fit(xi, mx, n), cmin = poly_fit(xi, mn, n);
        return {poly_eval(cmin, inventory), poly_eval(cmax, inventory)};
      }
    }
    throw OracleError(LSMC_ERR_ARGUMENT, "Unknown constraint kind.");
  }

  // CmdtyStorage.GetInjectWithdrawRange (CS.cs:86-100)
  Range inject_withdraw_range(int p, double inventory) const {
    double mn = min_inventory(p);
    if (inventory < mn) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory is below minimum allowed value during period.");
    double mx = max_inventory(p);
    if (inventory > mx) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory above maximum allowed value during period.");
    if (p >= P) return {0.0, 0.0};
    return constraint_range(p, inventory);
  }

  // InjectWithdrawCostNpv (LSV.cs:765-774) over the per-unit builder forms (CS.cs:322-409)
  double inject_withdraw_cost_npv(int p, double decision_volume) const {
    if (decision_volume > 0.0) return (s.injection_cost[p] * decision_volume) * s.injection_cost_df[p];
    return (s.withdrawal_cost[p] * std::fabs(-decision_volume)) * s.withdrawal_cost_df[p];
  }
  // CmdtyVolumeConsumedOnDecision (LSV.cs:757-763) over CS.cs:450-502
  double cmdty_consumed(int p, double decision_volume) const {
    if (decision_volume > 0.0) return s.consumed_on_inject[p] * std::fabs(decision_volume);
    return s.consumed_on_withdraw[p] * std::fabs(-decision_volume);
  }
  // Sum of CmdtyInventoryCost cash flows discounted (LSV.cs:247-248; CS.cs:546-568)
  double inventory_cost_npv(int p, double inventory) const { return (inventory * s.inventory_cost[p]) * s.inventory_cost_df[p]; }
  // TerminalStorageNpv (CS.cs:154-157)
  double terminal_npv(double price, double inventory) const {
    const double *c = s.terminal_npv_coeff;
    return ((c[3] * (price * inventory) + c[1] * price) + c[2] * inventory) + c[0];
  }

  // CmdtyStorage.InventorySpaceUpperBound / LowerBound (CS.cs:132-152) dispatching to the constraint
  double space_upper(int p, double next_lo, double next_hi) const {
    double cur_min = min_inventory(p), cur_max = max_inventory(p), loss = loss_pct(p);
    const double *xi = rinv(p), *mn = rmin(p), *mx = rmax(p);
    int n = rcount(p);
    switch (s.constraint_kind[p]) {
      case LSMC_CONSTRAINT_CONSTANT: {  // Constant…cs:50-56
        double solved = (next_hi - mn[0]) / (1 - loss);
        return std::min(solved, cur_max);
      }
      case LSMC_CONSTRAINT_PIECEWISE_LINEAR: {  // PiecewiseLinear…cs:74-116
        Range r = constraint_range(p, cur_max);
        double next_max_from_max = cur_max * (1 - loss) + r.max_rate;
        double next_min_from_max = cur_max * (1 - loss) + r.min_rate;
        if (next_min_from_max <= next_hi && next_lo <= next_max_from_max) return cur_max;
        double bracket_upper_inv = xi[n - 1];
        double bracket_upper_after = next_min_from_max;
        for (int i = n - 2; i >= 0; i--) {
          double bracket_lower_inv = xi[i];
          double bracket_lower_after = bracket_lower_inv * (1 - loss) + mn[i];
          if (bracket_lower_after <= next_hi && next_hi <= bracket_upper_after)
            return interpolate_linear_and_solve(bracket_lower_inv, bracket_lower_after, bracket_upper_inv, bracket_upper_after, next_hi);
          bracket_upper_after = bracket_lower_after;
          bracket_upper_inv = bracket_lower_inv;
        }
        throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
      }
      case LSMC_CONSTRAINT_STEP: {  // Step…cs:81-123
        Range r = constraint_range(p, cur_max);
        double next_max_from_max = cur_max * (1 - loss) + r.max_rate;
        double next_min_from_max = cur_max * (1 - loss) + r.min_rate;
        if (next_min_from_max <= next_hi && next_lo <= next_max_from_max) return cur_max;
        bool found = false;
        double result = 0.0;
        for (int i = 0; i < n - 1; i++) {
          double max_withdraw = mn[i];
          double lower_inv = xi[i], lower_after = lower_inv * (1 - loss) + max_withdraw;
          double upper_inv = xi[i + 1], upper_after = upper_inv * (1 - loss) + max_withdraw;
          if (lower_after <= next_hi && next_hi <= upper_after) {
            result = interpolate_linear_and_solve(lower_inv, lower_after, upper_inv, upper_after, next_hi);
            found = true;
          }
        }
        if (!found) throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
        return result;
      }
      case LSMC_CONSTRAINT_POLYNOMIAL: {  // Polynomial…cs:87-120 (note: no loss factor in the reachability test)
        std::vector<double> cmax = poly_fit(xi, mx, n), cmin = poly_fit(xi, mn, n), dmin = poly_diff(cmin);
        double next_max_from_max = cur_max + poly_eval(cmax, cur_max);
        double next_min_from_max = cur_max + poly_eval(cmin, cur_max);
        if (next_min_from_max <= next_hi && next_lo <= next_max_from_max) return cur_max;
        auto f = [&](double inv) { return inv * (1 - loss) + poly_eval(cmin, inv) - next_hi; };
        auto df = [&](double inv) { return (1 - loss) + poly_eval(dmin, inv); };
        double root;
        if (!robust_newton(f, df, cur_min, cur_max, s.newton_raphson_accuracy, s.newton_raphson_max_iterations,
                           s.newton_raphson_subdivision, &root))
          throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Cannot solve for the current period inventory space upper bound.");
        if (root < cur_min) throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Inventory constraints cannot be satisfied.");
        return std::min(root, cur_max);
      }
    }
    throw OracleError(LSMC_ERR_ARGUMENT, "Unknown constraint kind.");
  }

  double space_lower(int p, double next_lo, double next_hi) const {
    double cur_min = min_inventory(p), cur_max = max_inventory(p), loss = loss_pct(p);
    const double *xi = rinv(p), *mn = rmin(p), *mx = rmax(p);
    int n = rcount(p);
    switch (s.constraint_kind[p]) {
      case LSMC_CONSTRAINT_CONSTANT: {  // Constant…cs:58-65
        double solved = (next_lo - mx[0]) / (1 - loss);
        return std::max(solved, cur_min);
      }
      case LSMC_CONSTRAINT_PIECEWISE_LINEAR: {  // PiecewiseLinear…cs:118-160
        Range r = constraint_range(p, cur_min);
        double next_max_from_min = cur_min * (1 - loss) + r.max_rate;
        double next_min_from_min = cur_min * (1 - loss) + r.min_rate;
        if (next_min_from_min <= next_hi && next_lo <= next_max_from_min) return cur_min;
        double bracket_lower_inv = xi[0];
        double bracket_lower_after = next_max_from_min;
        for (int i = 1; i < n; i++) {
          double bracket_upper_inv = xi[i];
          double bracket_upper_after = bracket_upper_inv * (1 - loss) + mx[i];
          if (bracket_lower_after <= next_lo && next_lo <= bracket_upper_after)
            return interpolate_linear_and_solve(bracket_lower_inv, bracket_lower_after, bracket_upper_inv, bracket_upper_after, next_lo);
          bracket_lower_after = bracket_upper_after;
          bracket_lower_inv = bracket_upper_inv;
        }
        throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
      }
      case LSMC_CONSTRAINT_STEP: {  // Step…cs:125-166
        Range r = constraint_range(p, cur_min);
        double next_max_from_min = cur_min * (1 - loss) + r.max_rate;
        double next_min_from_min = cur_min * (1 - loss) + r.min_rate;
        if (next_min_from_min <= next_hi && next_lo <= next_max_from_min) return cur_min;
        bool found = false;
        double result = 0.0;
        for (int i = n - 2; i >= 0; i--) {
          double max_inject = mx[i];
          double lower_inv = xi[i], lower_after = lower_inv * (1 - loss) + max_inject;
          double upper_inv = xi[i + 1], upper_after = upper_inv * (1 - loss) + max_inject;
          if (lower_after <= next_lo && next_lo <= upper_after) {
            result = interpolate_linear_and_solve(lower_inv, lower_after, upper_inv, upper_after, next_lo);
            found = true;
          }
        }
        if (!found) throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Storage inventory constraints cannot be satisfied.");
        return result;
      }
      case LSMC_CONSTRAINT_POLYNOMIAL: {  // Polynomial…cs:122-155
        std::vector<double> cmax = poly_fit(xi, mx, n), cmin = poly_fit(xi, mn, n), dmax = poly_diff(cmax);
        double next_max_from_min = cur_min + poly_eval(cmax, cur_min);
        double next_min_from_min = cur_min + poly_eval(cmin, cur_min);
        if (next_min_from_min <= next_hi && next_lo <= next_max_from_min) return cur_min;
        auto f = [&](double inv) { return inv * (1 - loss) + poly_eval(cmax, inv) - next_lo; };
        auto df = [&](double inv) { return (1 - loss) + poly_eval(dmax, inv); };
        double root;
        if (!robust_newton(f, df, cur_min, cur_max, s.newton_raphson_accuracy, s.newton_raphson_max_iterations,
                           s.newton_raphson_subdivision, &root))
          throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Cannot solve for the current period inventory space lower bound.");
        if (root > cur_max) throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Inventory constraints cannot be satisfied.");
        return std::max(root, cur_min);
      }
    }
    throw OracleError(LSMC_ERR_ARGUMENT, "Unknown constraint kind.");
  }
};

// StorageHelper.CalculateInventorySpace (SH.cs:39-107).  out[i] is the range of period i+1.
void calculate_inventory_space(const Storage &st, double starting_inventory, double *out_min, double *out_max) {
  int P = st.P;
  std::vector<double> fwd_max(P), fwd_min(P), back_max(P), back_min(P);
  double min_fwd = starting_inventory, max_fwd = starting_inventory;
  for (int i = 0; i < P; i++) {
    double loss_pct = st.loss_pct(i);
    double rate_min = st.inject_withdraw_range(i, min_fwd).min_rate;
    double loss_at_min = loss_pct * min_fwd;
    double storage_min = st.min_inventory(i + 1);
    min_fwd = std::max(min_fwd - loss_at_min + rate_min, storage_min);
    fwd_min[i] = min_fwd;
    double rate_max = st.inject_withdraw_range(i, max_fwd).max_rate;
    double loss_at_max = loss_pct * max_fwd;
    double storage_max = st.max_inventory(i + 1);
    max_fwd = std::min(max_fwd - loss_at_max + rate_max, storage_max);
    fwd_max[i] = max_fwd;
  }
  back_max[P - 1] = st.s.must_be_empty_at_end ? 0 : st.max_inventory(P);
  back_min[P - 1] = st.s.must_be_empty_at_end ? 0 : st.min_inventory(P);
  for (int i = P - 2; i >= 0; i--) {
    int period = i + 1;  // periodBackLoop: end-1 for i = P-2
    back_max[i] = st.space_upper(period, back_min[i + 1], back_max[i + 1]);
    back_min[i] = st.space_lower(period, back_min[i + 1], back_max[i + 1]);
  }
  for (int i = 0; i < P; i++) {
    double space_max = std::min(fwd_max[i], back_max[i]);
    double space_min = std::max(fwd_min[i], back_min[i]);
    if (space_min > space_max) throw OracleError(LSMC_ERR_INVENTORY_CONSTRAINTS, "Inventory constraints cannot be fulfilled.");
    out_min[i] = space_min;
    out_max[i] = space_max;
  }
}

// FixedSpacingStateSpaceGridCalc.GetGridPoints (StateSpaceGridCalc/FixedSpacingStateSpaceGridCalc.cs:45-63)
std::vector<double> grid_points(double lower, double upper, double spacing) {
  if (lower > upper) throw OracleError(LSMC_ERR_ARGUMENT, "stateSpaceLowerBound cannot be above stateSpaceUpperBound");
  std::vector<double> out;
  out.push_back(lower);
  if (lower < upper) {
    double grid_point = lower;
    do {
      grid_point += spacing;
      out.push_back(std::min(grid_point, upper));
    } while (grid_point < upper);
  }
  return out;
}

// StorageHelper.CalculateBangBangDecisionSet (SH.cs:109-204)
std::vector<double> bang_bang_decision_set(Range range, double current_inventory, double inventory_loss, double next_min,
                                           double next_max, double tol, int extra) {
  if (next_min > next_max) throw OracleError(LSMC_ERR_ARGUMENT, "nextStepMinInventory cannot be higher than nextStepMaxInventory.");
  if (extra < 0) throw OracleError(LSMC_ERR_ARGUMENT, "numExtraDecisions must be non-negative.");
  double after_loss = current_inventory - inventory_loss;
  double after_max_withdraw = range.min_rate + after_loss;
  double withdraw_rate;
  if (after_max_withdraw > next_max) {
    if (after_max_withdraw - next_max < tol)
      withdraw_rate = next_max - after_loss;
    else
      throw OracleError(LSMC_ERR_ARGUMENT, "Inventory constraints cannot be fulfilled. This could potentially be fixed by increasing the numerical tolerance.");
  } else if (after_max_withdraw > next_min) {
    withdraw_rate = range.min_rate;
  } else {
    withdraw_rate = next_min - after_loss;
  }
  double after_max_inject = range.max_rate + after_loss;
  double inject_rate;
  if (after_max_inject < next_min) {
    if (next_min - after_max_inject < tol)
      inject_rate = next_min - after_loss;
    else
      throw OracleError(LSMC_ERR_ARGUMENT, "Inventory constraints cannot be fulfilled. This could potentially be fixed by increasing the numerical tolerance.");
  } else if (after_max_inject < next_max) {
    inject_rate = range.max_rate;
  } else {
    inject_rate = next_max - after_loss;
  }
  auto extras = [](double mn, double mx, int n, double *res) {  // SH.cs:199-204
    double inc = (mx - mn) / (n + 1);
    for (int i = 0; i < n; i++) res[i] = (i + 1) * inc + mn;
  };
  std::vector<double> d;
  if (withdraw_rate >= 0.0 || inject_rate <= 0.0) {
    d.assign(extra + 2, 0.0);
    d[0] = withdraw_rate;
    d[extra + 1] = inject_rate;
    if (extra > 0) extras(withdraw_rate, inject_rate, extra, &d[1]);
  } else {
    d.assign(2 * extra + 3, 0.0);
    d[0] = withdraw_rate;
    d[2 * extra + 2] = inject_rate;
    if (extra > 0) {
      extras(withdraw_rate, 0, extra, &d[1]);
      extras(0, inject_rate, extra, &d[extra + 2]);
    }
  }
  return d;
}

inline bool equals_within_tol(double a, double b, double tol) { return std::fabs(a - b) <= tol; }  // SH.cs:316

// StorageHelper.BisectInventorySpace (SH.cs:280-314)
void bisect_inventory_space(const double *grid, int n, double inventory, double tol, int *lower, int *upper) {
  if (n == 1 && equals_within_tol(inventory, grid[0], tol)) {
    *lower = *upper = 0;
    return;
  }
  int lo = 0, hi = n - 1, top = hi;
  while (hi > lo) {
    int mid = (lo + hi) / 2;
    double inv_mid = grid[mid];
    if (equals_within_tol(inventory, inv_mid, tol)) {
      *lower = *upper = mid;
      return;
    }
    if (inv_mid > inventory) {
      hi = mid;
    } else {
      int mid1 = mid + 1;
      double inv_mid1 = grid[mid1];
      if (inventory <= inv_mid1) {
        *lower = mid;
        *upper = mid1;
        return;
      }
      if (equals_within_tol(inventory, inv_mid1, tol)) {
        *lower = *upper = mid1;
        return;
      }
      if (mid1 == top) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory is outside of inventoryGrid bounds.");
      lo = mid;
    }
  }
  throw OracleError(LSMC_ERR_ARGUMENT, "Inventory is outside of inventoryGrid bounds.");
}

// StorageHelper.MaxValueAndIndex (SH.cs:206-221): strict >, first maximum wins
inline int max_value_and_index(const double *a, int n, double *max_out) {
  double mx = a[0];
  int idx = 0;
  for (int i = 1; i < n; i++)
    if (a[i] > mx) {
      mx = a[i];
      idx = i;
    }
  *max_out = mx;
  return idx;
}

// BasisFunctions.Generic / SpotPricePower / Ones (BasisFunctions/BasisFunctions.cs:34-135) + PopulateDesignMatrix
// (LSV.cs:838-855).  Column-major N x K.  std::pow(x, int) plays Math.Pow (SURVEY hard part 8).
void populate_design_matrix(const int32_t *powers, int K, int F, const double *spot, const double *const *factors, int64_t N,
                            double *x_colmajor) {
  for (int k = 0; k < K; k++) {
    const int32_t *pw = powers + (size_t)k * (1 + F);
    double *col = x_colmajor + (size_t)k * N;
    bool any_factor = false;
    for (int i = 0; i < F; i++) any_factor |= pw[1 + i] > 0;
    if (!any_factor) {
      if (pw[0] == 0) {
        for (int64_t n = 0; n < N; n++) col[n] = 1.0;
      } else {
        for (int64_t n = 0; n < N; n++) col[n] = std::pow(spot[n], (double)pw[0]);
      }
      continue;
    }
    if (pw[0] == 0) {
      for (int64_t n = 0; n < N; n++) col[n] = 1.0;
    } else {
      for (int64_t n = 0; n < N; n++) col[n] = std::pow(spot[n], (double)pw[0]);
    }
    for (int i = 0; i < F; i++) {
      if (pw[1 + i] > 0) {
        const double *x = factors[i];
        for (int64_t n = 0; n < N; n++) col[n] *= std::pow(x[n], (double)pw[1 + i]);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Regression: continuation values C[g'][n] for every next-period grid point (LSV.cs:191-212)
// ------------------------------------------------------------------------------------------------
struct Regression {
  int K;
  int64_t N;
  int mode;
  // mode 0
  std::vector<double> pinv;  // K x N row-major
  // mode 1
  std::vector<dd> q;  // N x K row-major, orthonormal basis in double-double
};

void regression_prepare(Regression &r, const double *x, int K, int64_t N, int mode) {
  r.K = K;
  r.N = N;
  r.mode = mode;
  if (mode == 0) {
    // designMatrix.QR(QRMethod.Thin) (LSV.cs:193): Householder on a copy; thin Q formed explicitly.
    std::vector<double> a(x, x + (size_t)N * K);  // column-major
    std::vector<double> rmat((size_t)K * K, 0.0), tau(K, 0.0);
    std::vector<std::vector<double>> vs(K);
    for (int k = 0; k < K; k++) {
      double *col = &a[(size_t)k * N];
      double norm2 = 0.0;
      for (int64_t i = k; i < N; i++) norm2 += col[i] * col[i];
      double norm = std::sqrt(norm2);
      double alpha = col[k] > 0 ? -norm : norm;
      std::vector<double> v(N, 0.0);
      for (int64_t i = k; i < N; i++) v[i] = col[i];
      v[k] -= alpha;
      double vnorm2 = 0.0;
      for (int64_t i = k; i < N; i++) vnorm2 += v[i] * v[i];
      tau[k] = vnorm2 > 0 ? 2.0 / vnorm2 : 0.0;
      for (int j = k; j < K; j++) {
        double *cj = &a[(size_t)j * N];
        double dot = 0.0;
        for (int64_t i = k; i < N; i++) dot += v[i] * cj[i];
        double f = tau[k] * dot;
        for (int64_t i = k; i < N; i++) cj[i] -= f * v[i];
      }
      for (int j = k; j < K; j++) rmat[(size_t)k * K + j] = a[(size_t)j * N + k];
      vs[k] = std::move(v);
    }
    // thin Q = H_0 ... H_{K-1} [I_K; 0]
    std::vector<double> q((size_t)N * K, 0.0);  // column-major
    for (int k = 0; k < K; k++) q[(size_t)k * N + k] = 1.0;
    for (int k = K - 1; k >= 0; k--) {
      const std::vector<double> &v = vs[k];
      for (int j = 0; j < K; j++) {
        double *cj = &q[(size_t)j * N];
        double dot = 0.0;
        for (int64_t i = k; i < N; i++) dot += v[i] * cj[i];
        double f = tau[k] * dot;
        for (int64_t i = k; i < N; i++) cj[i] -= f * v[i];
      }
    }
    // rInverse = R.Inverse() (LSV.cs:194): back substitution on the upper triangle
    std::vector<double> rinv((size_t)K * K, 0.0);
    for (int j = 0; j < K; j++) {
      for (int i = j; i >= 0; i--) {
        double sum = (i == j) ? 1.0 : 0.0;
        for (int l = i + 1; l <= j; l++) sum -= rmat[(size_t)i * K + l] * rinv[(size_t)l * K + j];
        double diag = rmat[(size_t)i * K + i];
        if (diag == 0.0) throw OracleError(LSMC_ERR_NUMERICAL, "Design matrix is rank deficient.");
        rinv[(size_t)i * K + j] = sum / diag;
      }
    }
    // pseudoInverse = rInverse * qTranspose (LSV.cs:195-196): K x N
    r.pinv.assign((size_t)K * N, 0.0);
    for (int i = 0; i < K; i++)
      for (int64_t n = 0; n < N; n++) {
        double sum = 0.0;
        for (int l = 0; l < K; l++) sum += rinv[(size_t)i * K + l] * q[(size_t)l * N + n];
        r.pinv[(size_t)i * N + n] = sum;
      }
  } else {
    // exact mode: Gram matrix, Cholesky factor R and R^-1 in double-double; Q = X R^-1 kept in double-double.
    std::vector<dd> gram((size_t)K * K, dd{0, 0});
    for (int i = 0; i < K; i++)
      for (int j = i; j < K; j++) {
        const double *ci = x + (size_t)i * N, *cj = x + (size_t)j * N;
        dd acc{0, 0};
        for (int64_t n = 0; n < N; n++) acc = dd_add(acc, two_prod(ci[n], cj[n]));
        gram[(size_t)i * K + j] = gram[(size_t)j * K + i] = acc;
      }
    std::vector<dd> R((size_t)K * K, dd{0, 0});  // upper triangular, gram = R^T R
    for (int j = 0; j < K; j++) {
      for (int i = 0; i <= j; i++) {
        dd sum = gram[(size_t)i * K + j];
        for (int l = 0; l < i; l++) sum = dd_sub(sum, dd_mul(R[(size_t)l * K + i], R[(size_t)l * K + j]));
        if (i == j) {
          if (!(sum.hi > 0.0)) throw OracleError(LSMC_ERR_NUMERICAL, "Design matrix is rank deficient.");
          R[(size_t)i * K + j] = dd_sqrt(sum);
        } else {
          R[(size_t)i * K + j] = dd_div(sum, R[(size_t)i * K + i]);
        }
      }
    }
    std::vector<dd> rinv((size_t)K * K, dd{0, 0});
    for (int j = 0; j < K; j++)
      for (int i = j; i >= 0; i--) {
        dd sum = (i == j) ? dd{1.0, 0.0} : dd{0.0, 0.0};
        for (int l = i + 1; l <= j; l++) sum = dd_sub(sum, dd_mul(R[(size_t)i * K + l], rinv[(size_t)l * K + j]));
        rinv[(size_t)i * K + j] = dd_div(sum, R[(size_t)i * K + i]);
      }
    r.q.assign((size_t)N * K, dd{0, 0});
#pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < N; n++)
      for (int j = 0; j < K; j++) {
        dd acc{0, 0};
        for (int l = 0; l <= j; l++) acc = dd_add(acc, dd_mul_d(rinv[(size_t)l * K + j], x[(size_t)l * N + n]));
        r.q[(size_t)n * K + j] = acc;
      }
  }
}

// regressResults = pseudoInverse * y; estimatedContinuationValues = designMatrix * regressResults (LSV.cs:204-205)
void regression_fit(const Regression &r, const double *x, const double *y, double *fitted, double *beta_out) {
  int K = r.K;
  int64_t N = r.N;
  if (r.mode == 0) {
    double beta[LSMC_MAX_BASIS];
    for (int k = 0; k < K; k++) {
      const double *row = &r.pinv[(size_t)k * N];
      double sum = 0.0;
      for (int64_t n = 0; n < N; n++) sum += row[n] * y[n];
      beta[k] = sum;
    }
    for (int64_t n = 0; n < N; n++) {
      double sum = 0.0;
      for (int k = 0; k < K; k++) sum += x[(size_t)k * N + n] * beta[k];
      fitted[n] = sum;
    }
    if (beta_out) std::memcpy(beta_out, beta, sizeof(double) * K);
  } else {
    dd gamma[LSMC_MAX_BASIS];
    for (int k = 0; k < K; k++) gamma[k] = dd{0, 0};
    for (int64_t n = 0; n < N; n++)
      for (int k = 0; k < K; k++) gamma[k] = dd_add(gamma[k], dd_mul_d(r.q[(size_t)n * K + k], y[n]));
    for (int64_t n = 0; n < N; n++) {
      dd acc{0, 0};
      for (int k = 0; k < K; k++) acc = dd_add(acc, dd_mul(r.q[(size_t)n * K + k], gamma[k]));
      fitted[n] = acc.hi;
    }
    if (beta_out)
      for (int k = 0; k < K; k++) beta_out[k] = gamma[k].hi;
  }
}

double mean_value(const double *v, int64_t n, int mode) {
  if (mode == 0) {  // Enumerable.Average: sequential sum / count (LSV.cs:181,776-783)
    double s = 0.0;
    for (int64_t i = 0; i < n; i++) s += v[i];
    return s / n;
  }
  dd s{0, 0};
  for (int64_t i = 0; i < n; i++) s = dd_add(s, dd{v[i], 0.0});
  return dd_div(s, dd{(double)n, 0.0}).hi;
}

}  // namespace

// ================================================================================================
// extern "C" surface (ctypes)
// ================================================================================================
extern "C" {

const char *oracle_last_error() { return g_last_error.c_str(); }

#define ORACLE_TRY try {
#define ORACLE_CATCH                         \
  }                                          \
  catch (const OracleError &e) {             \
    g_last_error = e.what();                 \
    return e.code;                           \
  }                                          \
  catch (const std::exception &e) {          \
    g_last_error = e.what();                 \
    return LSMC_ERR_ARGUMENT;                \
  }                                          \
  return LSMC_OK;

int oracle_calculate_inventory_space(const lsmc_storage *storage, double starting_inventory, double *out_min, double *out_max) {
  ORACLE_TRY
  Storage st(*storage);
  calculate_inventory_space(st, starting_inventory, out_min, out_max);
  ORACLE_CATCH
}

int oracle_bang_bang_decision_set(double min_rate, double max_rate, double inventory, double inventory_loss, double next_min,
                                  double next_max, double tol, int extra, double *out, int *count) {
  ORACLE_TRY
  std::vector<double> d = bang_bang_decision_set({min_rate, max_rate}, inventory, inventory_loss, next_min, next_max, tol, extra);
  *count = (int)d.size();
  std::copy(d.begin(), d.end(), out);
  ORACLE_CATCH
}

int oracle_bisect_inventory_space(const double *grid, int n, double inventory, double tol, int *lower, int *upper) {
  ORACLE_TRY
  bisect_inventory_space(grid, n, inventory, tol, lower, upper);
  ORACLE_CATCH
}

int oracle_grid_points(double lower, double upper, double spacing, double *out, int capacity, int *count) {
  ORACLE_TRY
  std::vector<double> g = grid_points(lower, upper, spacing);
  *count = (int)g.size();
  for (int i = 0; i < std::min<int>(capacity, (int)g.size()); i++) out[i] = g[i];
  ORACLE_CATCH
}

int oracle_inject_withdraw_range(const lsmc_storage *storage, int period, double inventory, double *min_rate, double *max_rate) {
  ORACLE_TRY
  Storage st(*storage);
  Range r = st.inject_withdraw_range(period, inventory);
  *min_rate = r.min_rate;
  *max_rate = r.max_rate;
  ORACLE_CATCH
}

int oracle_constraint_space_bounds(const lsmc_storage *storage, int period, double next_lo, double next_hi, double *lower, double *upper) {
  ORACLE_TRY
  Storage st(*storage);
  *upper = st.space_upper(period, next_lo, next_hi);
  *lower = st.space_lower(period, next_lo, next_hi);
  ORACLE_CATCH
}

int oracle_max_value_and_index(const double *a, int n, double *max_out, int *index) {
  ORACLE_TRY
  *index = max_value_and_index(a, n, max_out);
  ORACLE_CATCH
}

int oracle_populate_design_matrix(const int32_t *powers, int K, int F, const double *spot, const double *const *factors, int64_t N,
                                  double *x_colmajor) {
  ORACLE_TRY
  populate_design_matrix(powers, K, F, spot, factors, N, x_colmajor);
  ORACLE_CATCH
}

// Regression of one y on one design matrix: fitted values + coefficients (beta for mode 0, gamma for mode 1).
int oracle_regress(const double *x_colmajor, int K, int64_t N, const double *y, int mode, double *fitted, double *coeffs) {
  ORACLE_TRY
  Regression r;
  regression_prepare(r, x_colmajor, K, N, mode);
  regression_fit(r, x_colmajor, y, fitted, coeffs);
  ORACLE_CATCH
}

struct oracle_options {
  int32_t regression_mode;  // 0 qr (double), 1 exact (double-double)
  int32_t num_threads;      // OpenMP threads (<=0: default)
};

struct oracle_panels {  // all optional (NULL = skip); [P+1][N] unless noted
  double *inventory, *inject_withdraw_volume, *cmdty_consumed, *inventory_loss, *net_volume, *pv;
  double *decision_index;  // [P][N]
  double *backward_value_first;  // [N] storage value by sim at period 0 (the vector averaged at LSV.cs:623)
};

// LsmcStorageValuation.Calculate (LSV.cs:57-657) on caller-supplied simulations (LVP.cs:198-216).
int oracle_lsmc_calculate(const lsmc_problem *pb, const double *reg_spot, const double *const *reg_factors, const double *val_spot,
                          const double *const *val_factors, const oracle_options *opt, lsmc_results *res, oracle_panels *panels) {
  ORACLE_TRY
  const int mode = opt ? opt->regression_mode : 0;
#ifdef _OPENMP
  if (opt && opt->num_threads > 0) omp_set_num_threads(opt->num_threads);
#endif
  if (pb->inventory < 0) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory cannot be negative.");  // LSV.cs:64-65
  Storage st(pb->storage);
  const int P = st.P;
  const int64_t N = pb->num_sims;
  const int K = pb->num_basis, F = pb->num_factors;
  const int first_sim = pb->current_period_is_active ? 1 : 0;  // LVP.cs:177
  const double tol = pb->numerical_tolerance;
  const int extra = pb->extra_decisions;
  if (P < 1) throw OracleError(LSMC_ERR_ARGUMENT, "Storage has expired");

  auto spot_row = [&](const double *spot, int p) { return spot + (size_t)(p - first_sim) * N; };
  auto factor_rows = [&](const double *const *factors, int p, const double **out) {
    for (int i = 0; i < F; i++) out[i] = factors[i] + (size_t)(p - first_sim) * N;
  };

  // inventory space for periods 1..P (LSV.cs:91)
  std::vector<double> space_min(P + 1, 0.0), space_max(P + 1, 0.0);
  calculate_inventory_space(st, pb->inventory, &space_min[1], &space_max[1]);

  std::vector<std::vector<double>> grids(P + 1);
  grids[P] = grid_points(space_min[P], space_max[P], pb->grid_spacing);  // LSV.cs:111-114

  // terminal values (LSV.cs:116-131)
  std::vector<std::vector<double>> v_next(grids[P].size(), std::vector<double>(N));
  {
    const double *s_end = spot_row(reg_spot, P);
    for (size_t i = 0; i < grids[P].size(); i++)
      for (int64_t n = 0; n < N; n++) v_next[i][n] = st.terminal_npv(s_end[n], grids[P][i]);
  }

  std::vector<double> design((size_t)N * K);
  // regression coefficients per period/grid point are not returned by the reference; the forward pass
  // recomputes continuation values from them (LSV.cs:408-415).  mode 0 keeps beta, mode 1 keeps the fit objects.
  std::vector<std::vector<double>> beta_by_period(P);   // [p][g'*K + k]   (mode 0)
  std::vector<std::vector<dd>> gamma_by_period(P);      // [p][g'*K + k]   (mode 1)
  std::vector<std::vector<dd>> rinv_by_period(P);       // [p][K*K]        (mode 1)
  std::vector<double> current_period_cont;

  double progress = 0.0;
  (void)progress;

  // ---------------------------------------------------------------------------- backward induction (LSV.cs:169-346)
  for (int p = P - 1; p >= 0; p--) {
    const std::vector<double> &next_grid = grids[p + 1];
    const int Gn = (int)next_grid.size();
    std::vector<std::vector<double>> c_reg(Gn, std::vector<double>(N));
    const bool is_current = (p == 0 && pb->current_period_is_active);
    if (is_current) {  // LSV.cs:174-185
      current_period_cont.assign(Gn, 0.0);
      for (int i = 0; i < Gn; i++) {
        double expected = mean_value(v_next[i].data(), N, mode);
        std::fill(c_reg[i].begin(), c_reg[i].end(), expected);
        current_period_cont[i] = expected;
      }
    } else {  // LSV.cs:191-212
      const double *fac[LSMC_MAX_FACTORS];
      factor_rows(reg_factors, p, fac);
      populate_design_matrix(pb->basis_powers, K, F, spot_row(reg_spot, p), fac, N, design.data());
      if (mode == 0) {
        Regression r;
        regression_prepare(r, design.data(), K, N, 0);
        beta_by_period[p].assign((size_t)Gn * K, 0.0);
#pragma omp parallel for schedule(dynamic)
        for (int i = 0; i < Gn; i++)
          regression_fit(r, design.data(), v_next[i].data(), c_reg[i].data(), &beta_by_period[p][(size_t)i * K]);
      } else {
        // exact mode keeps R^-1 and gamma in double-double so the forward pass can evaluate the same fit on the
        // valuation simulations
        Regression r;
        regression_prepare(r, design.data(), K, N, 1);
        gamma_by_period[p].assign((size_t)Gn * K, dd{0, 0});
#pragma omp parallel for schedule(dynamic)
        for (int i = 0; i < Gn; i++) {
          dd gamma[LSMC_MAX_BASIS];
          for (int k = 0; k < K; k++) gamma[k] = dd{0, 0};
          const double *y = v_next[i].data();
          for (int64_t n = 0; n < N; n++)
            for (int k = 0; k < K; k++) gamma[k] = dd_add(gamma[k], dd_mul_d(r.q[(size_t)n * K + k], y[n]));
          for (int64_t n = 0; n < N; n++) {
            dd acc{0, 0};
            for (int k = 0; k < K; k++) acc = dd_add(acc, dd_mul(r.q[(size_t)n * K + k], gamma[k]));
            c_reg[i][n] = acc.hi;
          }
          for (int k = 0; k < K; k++) gamma_by_period[p][(size_t)i * K + k] = gamma[k];
        }
        // recover R^-1 in dd for the forward pass: recompute (cheap) from the Gram matrix
        std::vector<dd> gram((size_t)K * K, dd{0, 0}), R((size_t)K * K, dd{0, 0}), rinv((size_t)K * K, dd{0, 0});
        for (int i = 0; i < K; i++)
          for (int j = i; j < K; j++) {
            const double *ci = &design[(size_t)i * N], *cj = &design[(size_t)j * N];
            dd acc{0, 0};
            for (int64_t n = 0; n < N; n++) acc = dd_add(acc, two_prod(ci[n], cj[n]));
            gram[(size_t)i * K + j] = gram[(size_t)j * K + i] = acc;
          }
        for (int j = 0; j < K; j++)
          for (int i = 0; i <= j; i++) {
            dd sum = gram[(size_t)i * K + j];
            for (int l = 0; l < i; l++) sum = dd_sub(sum, dd_mul(R[(size_t)l * K + i], R[(size_t)l * K + j]));
            R[(size_t)i * K + j] = (i == j) ? dd_sqrt(sum) : dd_div(sum, R[(size_t)i * K + i]);
          }
        for (int j = 0; j < K; j++)
          for (int i = j; i >= 0; i--) {
            dd sum = (i == j) ? dd{1.0, 0.0} : dd{0.0, 0.0};
            for (int l = i + 1; l <= j; l++) sum = dd_sub(sum, dd_mul(R[(size_t)i * K + l], rinv[(size_t)l * K + j]));
            rinv[(size_t)i * K + j] = dd_div(sum, R[(size_t)i * K + i]);
          }
        rinv_by_period[p] = rinv;
      }
    }

    std::vector<double> grid;  // LSV.cs:215-223
    if (p == 0)
      grid = {pb->inventory};
    else
      grid = grid_points(space_min[p], space_max[p], pb->grid_spacing);
    const double next_min = space_min[p + 1], next_max = space_max[p + 1];  // LSV.cs:224
    const double df = pb->settle_df[p];                                       // LSV.cs:228-229
    std::vector<double> const_prices;
    const double *prices;
    if (is_current) {  // LSV.cs:232-236
      const_prices.assign(N, pb->forward_price[p]);
      prices = const_prices.data();
    } else {
      prices = spot_row(reg_spot, p);
    }

    const int G = (int)grid.size();
    std::vector<std::vector<double>> v_this(G, std::vector<double>(N));
    std::string loop_error;
    int loop_code = 0;
#pragma omp parallel for schedule(dynamic)
    for (int gi = 0; gi < G; gi++) {
      try {
        double inventory = grid[gi];
        Range range = st.inject_withdraw_range(p, inventory);                                       // LSV.cs:243
        double inventory_loss = st.loss_pct(p) * inventory;                                          // LSV.cs:244
        std::vector<double> decisions = bang_bang_decision_set(range, inventory, inventory_loss, next_min, next_max, tol, extra);
        double inventory_cost_npv = st.inventory_cost_npv(p, inventory);                             // LSV.cs:247-248
        const int D = (int)decisions.size();
        std::vector<double> cost_npv(D), consumed(D);
        std::vector<std::vector<double>> own_reg(D), own_act(D);
        std::vector<const double *> reg_by_d(D, nullptr), act_by_d(D, nullptr);
        for (int d = 0; d < D; d++) {
          double v = decisions[d];
          cost_npv[d] = st.inject_withdraw_cost_npv(p, v);  // LSV.cs:260
          consumed[d] = st.cmdty_consumed(p, v);            // LSV.cs:263
          double after = inventory + v - inventory_loss;    // LSV.cs:266
          for (int j = 0; j < Gn; j++) {                    // LSV.cs:267-305 (linear scan)
            double next_inv = next_grid[j];
            if (std::fabs(next_inv - after) < 1E-8) {  // hard-coded tolerance, LSV.cs:270
              reg_by_d[d] = c_reg[j].data();
              act_by_d[d] = v_next[j].data();
              break;
            }
            if (next_inv > after) {
              if (j == 0) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory after decision is below the next period inventory grid.");
              double lower_inv = next_grid[j - 1], upper_inv = next_inv;
              double space = upper_inv - lower_inv;
              double lower_w = (upper_inv - after) / space;
              double upper_w = 1.0 - lower_w;
              own_reg[d].resize(N);
              own_act[d].resize(N);
              const double *lr = c_reg[j - 1].data(), *ur = c_reg[j].data();
              const double *la = v_next[j - 1].data(), *ua = v_next[j].data();
              // WeightedAverage (LSV.cs:828-836): v1*w1 and v2*w2 rounded separately, then added
              for (int64_t n = 0; n < N; n++) {
                double a = lr[n] * lower_w, b = ur[n] * upper_w;
                own_reg[d][n] = b + a;
                double c = la[n] * lower_w, e = ua[n] * upper_w;
                own_act[d][n] = e + c;
              }
              reg_by_d[d] = own_reg[d].data();
              act_by_d[d] = own_act[d].data();
              break;
            }
          }
          if (!reg_by_d[d]) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory after decision is above the next period inventory grid.");
        }
        double npvs[LSMC_MAX_DECISIONS];
        double *out = v_this[gi].data();
        for (int64_t n = 0; n < N; n++) {  // LSV.cs:310-336
          double price = prices[n];
          for (int d = 0; d < D; d++) {
            double v = decisions[d];
            double inject_withdraw_npv = -v * price * df;
            double consumed_npv = -consumed[d] * price * df;
            double immediate = inject_withdraw_npv - cost_npv[d] + consumed_npv;
            double continuation = reg_by_d[d][n];
            npvs[d] = immediate + continuation - inventory_cost_npv;
          }
          double best;
          int idx = max_value_and_index(npvs, D, &best);
          double adjust = -reg_by_d[idx][n] + act_by_d[idx][n];
          out[n] = best + adjust;
        }
      } catch (const OracleError &e) {
#pragma omp critical
        {
          loop_error = e.what();
          loop_code = e.code;
        }
      }
    }
    if (loop_code) throw OracleError(loop_code, loop_error);
    grids[p] = grid;
    v_next.swap(v_this);
  }

  // ---------------------------------------------------------------------------- forward simulation (LSV.cs:356-612)
  int64_t snapped_total = 0;  // (path, period) pairs moved onto an inventory limit (the documented departure below)
  std::vector<double> pv_by_sim(N, 0.0), deltas(P + 1, 0.0), profile((size_t)(P + 1) * 5, 0.0);
  std::vector<double> inv_this(N, pb->inventory), inv_next(N, 0.0);
  std::vector<lsmc_trigger_prices> trig(P);
  std::memset(trig.data(), 0, sizeof(lsmc_trigger_prices) * P);
  const int num_trigger = LSMC_NUM_TRIGGER_VOLUMES;

  std::vector<double> vol_sim(N), cons_sim(N), loss_sim(N), pv_sim(N);
  std::vector<int> idx_sim(N);

  for (int p = 0; p < P; p++) {
    const std::vector<double> &next_grid = grids[p + 1];
    const int Gn = (int)next_grid.size();
    const bool is_current = (p == 0 && pb->current_period_is_active);
    std::vector<std::vector<double>> c_reg(Gn, std::vector<double>(N));
    if (is_current) {  // LSV.cs:397-405
      for (int i = 0; i < Gn; i++) std::fill(c_reg[i].begin(), c_reg[i].end(), current_period_cont[i]);
    } else {  // LSV.cs:408-415
      const double *fac[LSMC_MAX_FACTORS];
      factor_rows(val_factors, p, fac);
      populate_design_matrix(pb->basis_powers, K, F, spot_row(val_spot, p), fac, N, design.data());
      if (mode == 0) {
#pragma omp parallel for schedule(dynamic)
        for (int i = 0; i < Gn; i++) {
          const double *beta = &beta_by_period[p][(size_t)i * K];
          for (int64_t n = 0; n < N; n++) {
            double sum = 0.0;
            for (int k = 0; k < K; k++) sum += design[(size_t)k * N + n] * beta[k];
            c_reg[i][n] = sum;
          }
        }
      } else {
        const std::vector<dd> &rinv = rinv_by_period[p];
        std::vector<dd> q((size_t)N * K);
#pragma omp parallel for schedule(static)
        for (int64_t n = 0; n < N; n++)
          for (int j = 0; j < K; j++) {
            dd acc{0, 0};
            for (int l = 0; l <= j; l++) acc = dd_add(acc, dd_mul_d(rinv[(size_t)l * K + j], design[(size_t)l * N + n]));
            q[(size_t)n * K + j] = acc;
          }
#pragma omp parallel for schedule(dynamic)
        for (int i = 0; i < Gn; i++) {
          const dd *gamma = &gamma_by_period[p][(size_t)i * K];
          for (int64_t n = 0; n < N; n++) {
            dd acc{0, 0};
            for (int k = 0; k < K; k++) acc = dd_add(acc, dd_mul(q[(size_t)n * K + k], gamma[k]));
            c_reg[i][n] = acc.hi;
          }
        }
      }
    }
    const double df = pb->settle_df[p];
    const double df_delta = pb->discount_deltas ? df : 1.0;  // LSV.cs:420
    std::vector<double> const_prices;
    const double *prices;
    if (is_current) {
      const_prices.assign(N, pb->forward_price[p]);
      prices = const_prices.data();
    } else {
      prices = spot_row(val_spot, p);
    }
    const double next_min = space_min[p + 1], next_max = space_max[p + 1];  // LSV.cs:432

    std::string loop_error;
    int loop_code = 0;
    int64_t snapped_here = 0;
#pragma omp parallel for schedule(static) reduction(+ : snapped_here)
    for (int64_t n = 0; n < N; n++) {  // LSV.cs:445-508
      try {
        double price = prices[n];
        double inventory = inv_this[n];
        // DEPARTURE from the reference (documented in DESIGN.md, "rounding at the inventory limits"): inv + v - loss can land
        // a few ulps outside [MinInventory, MaxInventory] when a decision was constrained to hit the limit and the
        // period has an inventory loss; CmdtyStorage.GetInjectWithdrawRange (CS.cs:88-94) would then throw.  A breach
        // smaller than NumericalTolerance is snapped onto the limit instead (same rule in the CUDA forward kernel).
        {
          double lo = st.min_inventory(p), hi = st.max_inventory(p);
          if (inventory < lo && lo - inventory <= tol) inventory = lo, snapped_here++;
          if (inventory > hi && inventory - hi <= tol) inventory = hi, snapped_here++;
          inv_this[n] = inventory;
        }
        Range range = st.inject_withdraw_range(p, inventory);
        double inventory_loss = st.loss_pct(p) * inventory;
        std::vector<double> decisions = bang_bang_decision_set(range, inventory, inventory_loss, next_min, next_max, tol, extra);
        double inventory_cost_npv = st.inventory_cost_npv(p, inventory);
        const int D = (int)decisions.size();
        double npvs[LSMC_MAX_DECISIONS], cons[LSMC_MAX_DECISIONS], imm[LSMC_MAX_DECISIONS];
        for (int d = 0; d < D; d++) {
          double v = decisions[d];
          double after = inventory + v - inventory_loss;
          double consumed = st.cmdty_consumed(p, v);
          double inject_withdraw_npv = -v * price * df;
          double consumed_npv = -consumed * price * df;
          double cost_npv = st.inject_withdraw_cost_npv(p, v);
          double immediate = inject_withdraw_npv - cost_npv + consumed_npv - inventory_cost_npv;  // LSV.cs:473
          // InterpolateContinuationValue (LSV.cs:807-826)
          int lo, hi;
          bisect_inventory_space(next_grid.data(), Gn, after, tol, &lo, &hi);
          double continuation;
          if (lo == hi) {
            continuation = c_reg[lo][n];
          } else {
            double lower_inv = next_grid[lo], upper_inv = next_grid[hi];
            double space = upper_inv - lower_inv;
            double lower_w = (upper_inv - after) / space;
            double upper_w = 1.0 - lower_w;
            continuation = c_reg[lo][n] * lower_w + c_reg[hi][n] * upper_w;
          }
          npvs[d] = immediate + continuation;
          cons[d] = consumed;
          imm[d] = immediate;
        }
        double best;
        int idx = max_value_and_index(npvs, D, &best);
        double v_opt = decisions[idx];
        inv_next[n] = inventory + v_opt - inventory_loss;  // LSV.cs:485-486
        vol_sim[n] = v_opt;
        cons_sim[n] = cons[idx];
        loss_sim[n] = inventory_loss;
        pv_sim[n] = imm[idx];
        idx_sim[n] = idx;
      } catch (const OracleError &e) {
#pragma omp critical
        {
          loop_error = e.what();
          loop_code = e.code;
        }
      }
    }
    if (loop_code) throw OracleError(loop_code, loop_error);
    snapped_total += snapped_here;

    // sequential accumulation in sim order, as the reference's single loop does (LSV.cs:490-507)
    double sum_spot_vol = 0.0, sum_vol = 0.0, sum_cons = 0.0, sum_loss = 0.0, sum_pv = 0.0;
    for (int64_t n = 0; n < N; n++) {
      sum_spot_vol += -(vol_sim[n] + cons_sim[n]) * prices[n];
      sum_vol += vol_sim[n];
      sum_cons += cons_sim[n];
      sum_loss += loss_sim[n];
      sum_pv += pv_sim[n];
      pv_by_sim[n] += pv_sim[n];
    }
    if (panels) {
      size_t off = (size_t)p * N;
      for (int64_t n = 0; n < N; n++) {
        if (panels->inventory) panels->inventory[off + n] = inv_this[n];
        if (panels->inject_withdraw_volume) panels->inject_withdraw_volume[off + n] = vol_sim[n];
        if (panels->cmdty_consumed) panels->cmdty_consumed[off + n] = cons_sim[n];
        if (panels->inventory_loss) panels->inventory_loss[off + n] = loss_sim[n];
        if (panels->net_volume) panels->net_volume[off + n] = -vol_sim[n] - cons_sim[n];
        if (panels->pv) panels->pv[off + n] = pv_sim[n];
        if (panels->decision_index) panels->decision_index[off + n] = (double)idx_sim[n];
      }
    }
    double expected_inventory = mean_value(inv_this.data(), N, 0);  // Average (LSV.cs:510,776-783)
    double *pr = &profile[(size_t)p * 5];
    pr[0] = expected_inventory;
    pr[1] = sum_vol / N;
    pr[2] = sum_cons / N;
    pr[3] = sum_loss / N;
    pr[4] = sum_pv / N;
    deltas[p] = (sum_spot_vol / pb->forward_price[p] / N) * df_delta;  // LSV.cs:517

    // ---- trigger prices (LSV.cs:523-592)
    {
      lsmc_trigger_prices &tp = trig[p];
      double exp_loss = st.loss_pct(p) * expected_inventory;
      Range exp_range = st.inject_withdraw_range(p, expected_inventory);
      std::vector<double> dset = bang_bang_decision_set(exp_range, expected_inventory, exp_loss, next_min, next_max, tol, extra);
      auto average_continuation = [&](double after) {  // AverageContinuationValue (LSV.cs:785-805)
        int lo, hi;
        bisect_inventory_space(next_grid.data(), Gn, after, tol, &lo, &hi);
        if (lo == hi) return mean_value(c_reg[lo].data(), N, 0);
        double lower_inv = next_grid[lo], upper_inv = next_grid[hi];
        double space = upper_inv - lower_inv;
        double lower_w = (upper_inv - after) / space;
        double upper_w = 1.0 - lower_w;
        double s = 0.0;
        for (int64_t n = 0; n < N; n++) s += c_reg[lo][n] * lower_w + c_reg[hi][n] * upper_w;
        return s / N;
      };
      auto calc_trigger = [&](double trigger_volume, double alt_cont, double alt_volume, double alt_cost, double alt_consumed) {
        double after = expected_inventory + trigger_volume - exp_loss;  // LSV.cs:710
        double cont = average_continuation(after);
        double cont_change = cont - alt_cont;
        double excess = trigger_volume - alt_volume;
        double cost_change = st.inject_withdraw_cost_npv(p, trigger_volume) - alt_cost;
        double consumed_change = st.cmdty_consumed(p, trigger_volume) - alt_consumed;
        return (cont_change - cost_change) / (df * (excess + consumed_change));  // LSV.cs:720-721
      };
      double max_inject = *std::max_element(dset.begin(), dset.end());
      if (max_inject > 0) {
        double alt = 0.0;
        bool have = false;
        for (double d : dset)
          if (d >= 0 && (!have || d < alt)) {
            alt = d;
            have = true;
          }
        if (max_inject > alt) {
          double alt_cont = average_continuation(expected_inventory + alt - exp_loss);  // CalcAlternatives LSV.cs:746-755
          double alt_cost = st.inject_withdraw_cost_npv(p, alt);
          double alt_cons = st.cmdty_consumed(p, alt);
          double inc = (max_inject - alt) / num_trigger;  // CalcInjectTriggerPriceVolumes LSV.cs:725-734
          double vols[LSMC_NUM_TRIGGER_VOLUMES];
          vols[num_trigger - 1] = max_inject;
          for (int i = 1; i < num_trigger; i++) vols[i - 1] = alt + i * inc;
          for (int i = 0; i < num_trigger; i++) {
            tp.inject_volume[i] = vols[i];
            tp.inject_price[i] = calc_trigger(vols[i], alt_cont, alt, alt_cost, alt_cons);
          }
          tp.num_inject = num_trigger;
          tp.has_inject_price = 1;
          tp.max_inject_trigger_price = tp.inject_price[num_trigger - 1];
          tp.max_inject_volume = max_inject;
        }
      }
      double max_withdraw = *std::min_element(dset.begin(), dset.end());
      if (max_withdraw < 0) {
        double alt = 0.0;
        bool have = false;
        for (double d : dset)
          if (d <= 0 && (!have || d > alt)) {
            alt = d;
            have = true;
          }
        if (max_withdraw < alt) {
          double alt_cont = average_continuation(expected_inventory + alt - exp_loss);
          double alt_cost = st.inject_withdraw_cost_npv(p, alt);
          double alt_cons = st.cmdty_consumed(p, alt);
          double inc = (alt - max_withdraw) / num_trigger;  // CalcWithdrawTriggerPriceVolumes LSV.cs:736-744
          double vols[LSMC_NUM_TRIGGER_VOLUMES];
          for (int i = 0; i < num_trigger; i++) vols[i] = max_withdraw + i * inc;
          for (int i = 0; i < num_trigger; i++) {  // iterated in reverse (LSV.cs:576)
            double v = vols[num_trigger - 1 - i];
            tp.withdraw_volume[i] = v;
            tp.withdraw_price[i] = calc_trigger(v, alt_cont, alt, alt_cost, alt_cons);
          }
          tp.num_withdraw = num_trigger;
          tp.has_withdraw_price = 1;
          tp.max_withdraw_trigger_price = tp.withdraw_price[0];  // LSV.cs:584
          tp.max_withdraw_volume = max_withdraw;
        }
      }
    }
    inv_this.swap(inv_next);
  }
  // after the loop inv_this holds the inventories at the end period ("nextPeriodInventories", LSV.cs:599)
  double end_period_pv = 0.0;
  if (!pb->storage.must_be_empty_at_end) {  // LSV.cs:596-612
    const double *s_end = spot_row(reg_spot, P);  // regression sims' end prices
    double terminal_pv = 0.0;
    for (int64_t n = 0; n < N; n++) {
      double t = st.terminal_npv(s_end[n], inv_this[n]);
      terminal_pv += t;
      double add = pb->reference_compat_terminal_pv ? terminal_pv : t;  // cumulative-sum quirk
      if (panels && panels->pv) panels->pv[(size_t)P * N + n] = add;
      pv_by_sim[n] += add;
    }
    end_period_pv = terminal_pv / N;
  }
  if (panels && panels->inventory)
    for (int64_t n = 0; n < N; n++) panels->inventory[(size_t)P * N + n] = inv_this[n];
  if (panels && panels->backward_value_first) std::memcpy(panels->backward_value_first, v_next[0].data(), sizeof(double) * N);

  double forward_npv = mean_value(pv_by_sim.data(), N, 0);  // LSV.cs:617
  // MathNet Statistics.StandardDeviation: sample (N-1) standard deviation, streaming update (LSV.cs:618)
  double variance = 0.0;
  if (N > 1) {
    double t = pv_by_sim[0];
    for (int64_t i = 1; i < N; i++) {
      t += pv_by_sim[i];
      double diff = ((double)(i + 1) * pv_by_sim[i]) - t;
      variance += (diff * diff) / ((i + 1.0) * i);
    }
    variance /= (double)(N - 1);
  }
  res->snapped_inventories = snapped_total;
  res->npv = forward_npv;
  res->std_error = std::sqrt(variance) / std::sqrt((double)N);
  res->backward_npv = mean_value(v_next[0].data(), N, 0);  // LSV.cs:623
  double *pr = &profile[(size_t)P * 5];                    // LSV.cs:627-629
  pr[0] = mean_value(inv_this.data(), N, 0);
  pr[4] = end_period_pv;
  if (res->deltas) std::copy(deltas.begin(), deltas.end(), res->deltas);
  if (res->profile) std::copy(profile.begin(), profile.end(), res->profile);
  if (res->trigger_prices) std::memcpy(res->trigger_prices, trig.data(), sizeof(lsmc_trigger_prices) * P);
  if (res->pv_by_sim) std::copy(pv_by_sim.begin(), pv_by_sim.end(), res->pv_by_sim);
  ORACLE_CATCH
}

// Grids used by oracle_lsmc_calculate, for tests: returns count for `period`.
int oracle_period_grid(const lsmc_problem *pb, int period, double *out, int capacity, int *count) {
  ORACLE_TRY
  Storage st(pb->storage);
  int P = st.P;
  std::vector<double> space_min(P + 1, 0.0), space_max(P + 1, 0.0);
  calculate_inventory_space(st, pb->inventory, &space_min[1], &space_max[1]);
  std::vector<double> g;
  if (period == 0)
    g = {pb->inventory};
  else
    g = grid_points(space_min[period], space_max[period], pb->grid_spacing);
  *count = (int)g.size();
  for (int i = 0; i < std::min<int>(capacity, (int)g.size()); i++) out[i] = g[i];
  ORACLE_CATCH
}

// IntrinsicStorageValuation.Calculate (IntrinsicValuation/IntrinsicStorageValuation.cs:120-322) with the linear inventory-space
// interpolator (InterpolatorFactories/LinearInterpolatorFactory.cs:36-49 -> MathNet LinearSpline, restated in linear_spline
// above) and the fixed-spacing grid: deterministic dynamic programme on the forward curve.  It shares the inventory space,
// grid, bang-bang decision set and MaxValueAndIndex helpers with the LSMC valuation, and the reference holds exact golden
// values for it (src/Cmdty.Storage.Python/tests/test_multi_factor.py:102,123), which is what pins those helpers end to end.
// profile is [P+1][5]: inventory after the decision, inject/withdraw volume, commodity consumed, inventory loss, period PV.
int oracle_intrinsic_calculate(const lsmc_problem *pb, double *npv_out, double *profile) {
  ORACLE_TRY
  if (pb->inventory < 0) throw OracleError(LSMC_ERR_ARGUMENT, "Inventory cannot be negative.");  // :127-128
  Storage st(pb->storage);
  const int P = st.P;
  const double tol = pb->numerical_tolerance;
  if (P < 1) throw OracleError(LSMC_ERR_ARGUMENT, "Storage has expired");
  std::vector<double> space_min(P + 1, 0.0), space_max(P + 1, 0.0);
  calculate_inventory_space(st, pb->inventory, &space_min[1], &space_max[1]);  // :154

  // value functions of periods 1..P as interpolation tables (period P: the terminal function, :188-190)
  std::vector<std::vector<double>> grid(P + 1), value(P + 1);
  auto continuation = [&](int period, double inventory) -> double {
    if (period == P) return st.terminal_npv(pb->forward_price[P], inventory);
    const std::vector<double> &x = grid[period], &y = value[period];
    if (x.size() == 1) return y[0];  // LinearInterpolatorFactory.cs:43-47
    return linear_spline(x.data(), y.data(), (int)x.size(), inventory);
  };
  struct Decision {
    double storage_npv, inject_withdraw, consumed, loss, period_pv;
  };
  // OptimalDecisionAndValue (:260-291) + StorageValueForDecision (:294-322)
  auto optimal = [&](int p, double inventory) -> Decision {
    const Range range = st.inject_withdraw_range(p, inventory);
    const double loss = st.loss_pct(p) * inventory;
    const std::vector<double> decisions = bang_bang_decision_set(range, inventory, loss, space_min[p + 1], space_max[p + 1], tol, 0);
    const double price = pb->forward_price[p], df = pb->settle_df[p];
    std::vector<double> values(decisions.size()), consumed(decisions.size()), pvs(decisions.size());
    for (size_t j = 0; j < decisions.size(); j++) {
      const double v = decisions[j];
      const double after = inventory + v - loss;
      const double cont = continuation(p + 1, after);
      const double inject_withdraw_npv = -v * price * df;
      const double inventory_cost_npv = st.inventory_cost_npv(p, inventory);
      const double decision_cost_npv = st.inject_withdraw_cost_npv(p, v);
      const double used = st.cmdty_consumed(p, v);
      const double used_npv = -used * price * df;
      const double period_pv = inject_withdraw_npv - decision_cost_npv + used_npv - inventory_cost_npv;
      values[j] = cont + period_pv;
      consumed[j] = used;
      pvs[j] = period_pv;
    }
    double best;
    const int idx = max_value_and_index(values.data(), (int)values.size(), &best);
    return {best, decisions[idx], consumed[idx], loss, pvs[idx]};
  };
  for (int p = P - 1; p >= 1; p--) {  // :196-223
    grid[p] = grid_points(space_min[p], space_max[p], pb->grid_spacing);
    value[p].resize(grid[p].size());
    for (size_t i = 0; i < grid[p].size(); i++) value[p][i] = optimal(p, grid[p][i]).storage_npv;
  }
  // forward from the starting inventory (:225-255)
  double inventory = pb->inventory, npv = 0.0;
  for (int p = 0; p <= P; p++) {
    double *row = profile + (size_t)p * 5;
    if (p == P) {
      const double end_npv = pb->storage.must_be_empty_at_end ? 0.0 : st.terminal_npv(pb->forward_price[P], inventory);
      row[0] = inventory, row[1] = 0.0, row[2] = 0.0, row[3] = 0.0, row[4] = end_npv;
    } else {
      const Decision d = optimal(p, inventory);
      inventory += d.inject_withdraw - d.loss;
      row[0] = inventory, row[1] = d.inject_withdraw, row[2] = d.consumed, row[3] = d.loss, row[4] = d.period_pv;
    }
    npv += row[4];  // storageProfiles.Sum(profile => profile.PeriodPv), :257
  }
  *npv_out = npv;
  ORACLE_CATCH
}

int oracle_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
