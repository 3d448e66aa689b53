"""Host-side mirror of the reference's storage entity: CmdtyStorage<T>.Builder (StorageEntity/CmdtyStorage.cs:171-569)
and the ratchet extensions (StorageEntity/CmdtyStorageBuilderExtensions.cs:38-257).

Same method names, argument meaning and error behaviour; the difference is that the builder records *structured*
parameters (per-period numbers and ratchet tables) instead of opaque Func<>s, because the device evaluates the storage
at per-path inventories (LsmcStorageValuation.cs:450-471).  Overloads that take arbitrary Func<period, inventory, volume>
(CmdtyStorage.cs:371,411,473,504,539) are rejected with UnsupportedStorageFeature.
"""
import ctypes as C

import numpy as np

from . import _abi
from .errors import ArgumentException, UnsupportedStorageFeature


class InjectWithdrawRange:
    """InjectWithdrawRange.cs: (MinInjectWithdrawRate, MaxInjectWithdrawRate)."""
    __slots__ = ("MinInjectWithdrawRate", "MaxInjectWithdrawRate")

    def __init__(self, min_rate, max_rate):
        self.MinInjectWithdrawRate = float(min_rate)
        self.MaxInjectWithdrawRate = float(max_rate)


class _Constraint:
    """Structured IInjectWithdrawConstraint: kind + rows (inventory, min_rate, max_rate) sorted by inventory."""

    def __init__(self, kind, rows, newton=(1e-10, 100, 20)):
        self.kind = kind
        self.rows = [(float(i), float(lo), float(hi)) for (i, lo, hi) in rows]
        self.newton = newton


def ConstantInjectWithdrawConstraint(min_rate, max_rate):
    return _Constraint(_abi.LSMC_CONSTRAINT_CONSTANT, [(0.0, min_rate, max_rate)])


def _rows_from(ranges):
    rows = []
    for r in ranges:
        inventory, rng = r
        if isinstance(rng, InjectWithdrawRange):
            rows.append((inventory, rng.MinInjectWithdrawRate, rng.MaxInjectWithdrawRate))
        else:
            rows.append((inventory, rng[0], rng[1]))
    return rows


def PiecewiseLinearInjectWithdrawConstraint(ranges):
    rows = sorted(_rows_from(ranges), key=lambda r: r[0])  # PiecewiseLinear…cs:45
    if len(rows) < 2:
        raise ArgumentException("Inject/withdraw ranges collection must contain at least two elements.")
    return _Constraint(_abi.LSMC_CONSTRAINT_PIECEWISE_LINEAR, rows)


def StepInjectWithdrawConstraint(ranges):
    rows = sorted(_rows_from(ranges), key=lambda r: r[0])  # Step…cs:42-68
    if len(rows) < 2:
        raise ArgumentException("Inject/withdraw ranges collection must contain at least two elements.")
    if abs(rows[-2][2] - rows[-1][2]) > 1e-12:
        raise ArgumentException("Top two ratchets do not have he same max injection rate.")
    if abs(rows[-2][1] - rows[-1][1]) > 1e-12:
        raise ArgumentException("Top two ratchets do not have he same max withdrawal rate.")
    for i in range(1, len(rows) - 1):
        if rows[i][2] > rows[i - 1][2]:
            raise ArgumentException("Ratchet injection rates cannot increase with inventory.")
        if rows[i][1] > rows[i - 1][1]:
            raise ArgumentException("Ratchet withdrawal rates cannot decrease with inventory.")
    return _Constraint(_abi.LSMC_CONSTRAINT_STEP, rows)


def PolynomialInjectWithdrawConstraint(ranges, newtonRaphsonAccuracy=1e-10, newtonRaphsonMaxNumIterations=100,
                                       newtonRaphsonSubdivision=20):
    if newtonRaphsonAccuracy < 0:
        raise ArgumentException("Newton Raphson Accuracy must be positive number.")
    if newtonRaphsonMaxNumIterations < 2:
        raise ArgumentException("Newton Raphson max number of iteration must be at least 2.")
    if newtonRaphsonSubdivision < 1:
        raise ArgumentException("Newton Raphson subdivision must be at least 1.")
    rows = _rows_from(ranges)  # Polynomial…cs:58-64 keeps the caller's order
    if len(rows) < 2:
        raise ArgumentException("At least 2 inject/withdraw constraints must be specified")
    return _Constraint(_abi.LSMC_CONSTRAINT_POLYNOMIAL, rows,
                       (newtonRaphsonAccuracy, newtonRaphsonMaxNumIterations, newtonRaphsonSubdivision))


class CmdtyStorage:
    """Immutable storage description; see CmdtyStorage<T> (CmdtyStorage.cs:39-169)."""

    def __init__(self, b):
        self.StartPeriod = b._start
        self.EndPeriod = b._end
        self.MustBeEmptyAtEnd = b._must_be_empty
        self._constraint = b._constraint        # period -> _Constraint
        self._min_inventory = b._min_inventory  # period -> float
        self._max_inventory_raw = b._max_inventory
        self._inject_cost = b._inject_cost      # period -> (per-unit cost, cash-flow Day)
        self._withdraw_cost = b._withdraw_cost
        self._inject_consumed = b._inject_consumed  # period -> fraction
        self._withdraw_consumed = b._withdraw_consumed
        self._inventory_loss = b._inventory_loss
        self._inventory_cost = b._inventory_cost  # period -> per-unit cost (cash flow on period.First<Day>())
        self._terminal = b._terminal              # (c0, c1, c2, c3)

    Builder = None  # set below

    def MinInventory(self, period):
        return self._min_inventory(period)

    def MaxInventory(self, period):  # CmdtyStorage.cs:435-438
        if self.MustBeEmptyAtEnd and period >= self.EndPeriod:
            return 0.0
        return self._max_inventory_raw(period)

    def TerminalStorageNpv(self, cmdtyPrice, finalInventory):
        c = self._terminal
        return ((c[3] * (cmdtyPrice * finalInventory) + c[1] * cmdtyPrice) + c[2] * finalInventory) + c[0]

    def CmdtyInventoryPercentLoss(self, period):
        return self._inventory_loss(period)

    def GetInjectWithdrawRange(self, period, inventory):
        """CmdtyStorage.cs:86-100, evaluated by the native library's host code (no GPU needed)."""
        flat = self.flatten(period, lambda d: 1.0)
        mn, mx = C.c_double(), C.c_double()
        lib = _abi.load_library()
        from .errors import check_status
        check_status(lib.lsmc_inject_withdraw_range(C.byref(flat.struct), 0, float(inventory), C.byref(mn), C.byref(mx)), None)
        return InjectWithdrawRange(mn.value, mx.value)

    def flatten(self, start_active, discount_to_current_day):
        """Builds the lsmc_storage descriptor for periods start_active..EndPeriod (index 0..P)."""
        return _FlatStorage(self, start_active, discount_to_current_day)


class _FlatStorage:
    def __init__(self, st, start_active, df):
        P = st.EndPeriod.OffsetFrom(start_active)
        if P < 0:
            raise ArgumentException("Storage has expired")
        periods = [start_active.Offset(i) for i in range(P + 1)]
        a = {}
        a["min_inventory"] = np.array([st._min_inventory(p) for p in periods], dtype=np.float64)
        a["max_inventory"] = np.array([st._max_inventory_raw(p) for p in periods], dtype=np.float64)
        kinds, begins, counts, rows, seen = [], [], [], [], {}
        newton = (1e-10, 100, 20)
        for p in periods[:P]:
            c = st._constraint(p)
            key = id(c)
            if key not in seen:
                seen[key] = (len(rows), len(c.rows))
                rows.extend(c.rows)
            b, n = seen[key]
            kinds.append(c.kind)
            begins.append(b)
            counts.append(n)
            if c.kind == _abi.LSMC_CONSTRAINT_POLYNOMIAL:
                newton = c.newton
        a["constraint_kind"] = np.array(kinds, dtype=np.int32)
        a["ratchet_begin"] = np.array(begins, dtype=np.int32)
        a["ratchet_count"] = np.array(counts, dtype=np.int32)
        rows = rows or [(0.0, 0.0, 0.0)]
        a["ratchet_inventory"] = np.array([r[0] for r in rows], dtype=np.float64)
        a["ratchet_min_rate"] = np.array([r[1] for r in rows], dtype=np.float64)
        a["ratchet_max_rate"] = np.array([r[2] for r in rows], dtype=np.float64)

        def cost_arrays(fn):
            cost, dfs = [], []
            for p in periods[:P]:
                per_unit, day = fn(p)
                cost.append(per_unit)
                dfs.append(df(day))
            return np.array(cost, dtype=np.float64), np.array(dfs, dtype=np.float64)

        a["injection_cost"], a["injection_cost_df"] = cost_arrays(st._inject_cost)
        a["withdrawal_cost"], a["withdrawal_cost_df"] = cost_arrays(st._withdraw_cost)
        a["consumed_on_inject"] = np.array([st._inject_consumed(p) for p in periods[:P]], dtype=np.float64)
        a["consumed_on_withdraw"] = np.array([st._withdraw_consumed(p) for p in periods[:P]], dtype=np.float64)
        a["inventory_loss"] = np.array([st._inventory_loss(p) for p in periods[:P]], dtype=np.float64)
        a["inventory_cost"] = np.array([st._inventory_cost(p) for p in periods[:P]], dtype=np.float64)
        a["inventory_cost_df"] = np.array([df(p.First()) for p in periods[:P]], dtype=np.float64)
        for k in list(a):
            if a[k].size == 0:  # keep pointers valid for P == 0
                a[k] = np.zeros(1, dtype=a[k].dtype)
        self.arrays = a
        s = _abi.lsmc_storage()
        s.num_periods = P
        s.must_be_empty_at_end = 1 if st.MustBeEmptyAtEnd else 0
        for name, arr in a.items():
            ptr_t = _abi.c_int32_p if arr.dtype == np.int32 else _abi.c_double_p
            setattr(s, name, arr.ctypes.data_as(ptr_t))
        s.num_ratchet_rows = len(rows)
        s.terminal_npv_coeff = (C.c_double * 4)(*st._terminal)
        s.newton_raphson_accuracy = newton[0]
        s.newton_raphson_max_iterations = newton[1]
        s.newton_raphson_subdivision = newton[2]
        self.struct = s
        self.num_periods = P
        self.periods = periods


def _series_lookup(series, what, start, end_needed):
    """TimeSeries<T,double> given as dict {period: value}; validates coverage like CmdtyStorage.cs:343-358."""
    if not series:
        raise ArgumentException(f"{what} time series cannot be empty.")
    keys = sorted(series)
    if keys[0] > start:
        raise ArgumentException(f"{what} time series starts at {keys[0]} which is later than the storage start period {start}.")
    if keys[-1] < end_needed:
        raise ArgumentException(f"{what} time series ends at {keys[-1]} which is earlier than {end_needed}.")
    return lambda p: float(series[p])


class _StorageBuilder:
    """CmdtyStorage<T>.Builder fluent interface (CmdtyStorage.cs:173-569)."""

    def __init__(self):
        self._start = self._end = None
        self._constraint = None
        self._min_inventory = self._max_inventory = None
        self._inject_cost = self._withdraw_cost = None
        self._inject_consumed = self._withdraw_consumed = None
        self._inventory_loss = None
        self._inventory_cost = None
        self._terminal = (0.0, 0.0, 0.0, 0.0)
        self._must_be_empty = False

    # -- active period (CmdtyStorage.cs:194-201)
    def WithActiveTimePeriod(self, start, end):
        if start >= end:
            raise ArgumentException("Storage start period must be before end period.")
        self._start, self._end = start, end
        return self

    # -- inject/withdraw constraints (CmdtyStorage.cs:203-245; CmdtyStorageBuilderExtensions.cs:38-140)
    def WithConstantInjectWithdrawRange(self, minInjectWithdrawRate, maxInjectWithdrawRate):
        c = ConstantInjectWithdrawConstraint(minInjectWithdrawRate, maxInjectWithdrawRate)
        self._constraint = lambda p: c
        return self

    def WithInjectWithdrawConstraint(self, constraint):
        if constraint is None:
            raise ArgumentException("injectWithdrawConstraint is null")
        if callable(constraint):
            self._constraint = constraint
        else:
            self._constraint = lambda p: constraint
        return self

    def WithInjectWithdrawRangeSeries(self, series):
        lookup_keys = sorted(series)
        if not lookup_keys:
            raise ArgumentException("Inject/withdraw range time series cannot be empty.")
        if lookup_keys[0] > self._start:
            raise ArgumentException("Inject/withdraw range time series starts later than the storage start period.")
        if lookup_keys[-1] < self._end:
            raise ArgumentException("Inject/withdraw range time series ends earlier than the storage end period.")
        cs = {p: ConstantInjectWithdrawConstraint(r.MinInjectWithdrawRate, r.MaxInjectWithdrawRate) for p, r in series.items()}
        self._constraint = lambda p: cs[p]
        return self

    def _add_ratchets(self, injectWithdrawRanges, factory):
        # AddInjectWithdrawRanges (CmdtyStorageBuilderExtensions.cs:142-257)
        items = []
        for period, ranges in injectWithdrawRanges:
            if period is None:
                raise ArgumentException("Null Period in collection.")
            if ranges is None:
                raise ArgumentException("Null InjectWithdrawRanges in collection.")
            rows = _rows_from(ranges)
            if len(rows) < 2:
                raise ArgumentException(f"Period {period} contains less than 2 inject/withdraw/inventory constraints.")
            if len(rows) == 2 and rows[0][1] == rows[1][1] and rows[0][2] == rows[1][2]:
                constraint = ConstantInjectWithdrawConstraint(rows[0][1], rows[0][2])
            else:
                constraint = factory(ranges)
            inv = [r[0] for r in rows]
            items.append((period, constraint, min(inv), max(inv)))
        if not items:
            raise ArgumentException("No inject/withdraw constrains provided.")
        items.sort(key=lambda it: it[0])
        for a, b in zip(items, items[1:]):
            if a[0] == b[0]:
                raise ArgumentException("Repeated periods found in inject/withdraw ranges.")

        def find(period):  # piecewise-flat in time; beyond the last key the last entry applies
            chosen = None
            for it in items:
                if it[0] <= period:
                    chosen = it
                else:
                    break
            if chosen is None:
                raise ArgumentException(f"No inject/withdraw constraint for period {period}.")
            return chosen

        self._constraint = lambda p: find(p)[1]
        self._min_inventory = lambda p: find(p)[2]
        self._max_inventory = lambda p: find(p)[3]
        return self

    def WithTimeAndInventoryVaryingInjectWithdrawRatesPiecewiseLinear(self, injectWithdrawRanges):
        return self._add_ratchets(injectWithdrawRanges, PiecewiseLinearInjectWithdrawConstraint)

    def WithStepRatchets(self, injectWithdrawRanges):
        return self._add_ratchets(injectWithdrawRanges, StepInjectWithdrawConstraint)

    def WithTimeAndInventoryVaryingInjectWithdrawRatesPolynomial(self, injectWithdrawRanges, newtonRaphsonAccuracy=1e-10,
                                                                 newtonRaphsonMaxNumIterations=100, newtonRaphsonSubdivision=20):
        return self._add_ratchets(
            injectWithdrawRanges,
            lambda r: PolynomialInjectWithdrawConstraint(r, newtonRaphsonAccuracy, newtonRaphsonMaxNumIterations,
                                                         newtonRaphsonSubdivision))

    # -- inventory limits (CmdtyStorage.cs:247-320)
    def WithZeroMinInventory(self):
        self._min_inventory = lambda p: 0.0
        return self

    def WithConstantMinInventory(self, minInventory):
        if minInventory < 0:
            raise ArgumentException("Minimum inventory must be non-negative.")
        self._min_inventory = lambda p: float(minInventory)
        return self

    def WithMinInventoryTimeSeries(self, series):
        self._min_inventory = _series_lookup(series, "Minimum inventory", self._start, self._end)
        return self

    def WithConstantMaxInventory(self, maxInventory):
        if maxInventory < 0:
            raise ArgumentException("Maximum inventory must be non-negative.")
        self._max_inventory = lambda p: float(maxInventory)
        return self

    def WithMaxInventoryTimeSeries(self, series):
        self._max_inventory = _series_lookup(series, "Maximum inventory", self._start, self._end)
        return self

    # -- injection / withdrawal cost (CmdtyStorage.cs:322-416)
    def WithPerUnitInjectionCost(self, perVolumeUnitCost, cashFlowDate=None):
        if perVolumeUnitCost < 0:
            raise ArgumentException("Per unit inject cost must be non-negative.")
        date_fn = cashFlowDate or (lambda p: p.First())
        self._inject_cost = lambda p: (float(perVolumeUnitCost), date_fn(p))
        return self

    def WithPerUnitInjectionCostTimeSeries(self, series):
        f = _series_lookup(series, "Per unit injection cost", self._start, self._end.Offset(-1))
        self._inject_cost = lambda p: (f(p), p.First())
        return self

    def WithInjectionCost(self, injectionCost):
        raise UnsupportedStorageFeature("WithInjectionCost(Func) has no device-evaluable form (CmdtyStorage.cs:371); "
                                        "use a per-unit cost.")

    def WithPerUnitWithdrawalCost(self, perVolumeUnitCost, cashFlowDate=None):
        if perVolumeUnitCost < 0:
            raise ArgumentException("Per unit withdrawal cost must be non-negative.")
        date_fn = cashFlowDate or (lambda p: p.First())
        self._withdraw_cost = lambda p: (float(perVolumeUnitCost), date_fn(p))
        return self

    def WithPerUnitWithdrawalCostTimeSeries(self, series):
        f = _series_lookup(series, "Per unit withdrawal cost", self._start, self._end.Offset(-1))
        self._withdraw_cost = lambda p: (f(p), p.First())
        return self

    def WithPerUnitWithdrawalCostByPeriod(self, costAndCashFlowDate):
        """Structured form of WithWithdrawalCost(Func) (CmdtyStorage.cs:411) for costs that are proportional to the withdrawn
        volume: costAndCashFlowDate(period) -> (cost per unit withdrawn in that period, day the cash flow is paid).  This is what
        the reference's own call-option fixture passes as a Func (tests/Cmdty.Storage.Test/TestHelper.cs:177-195: a strike per
        withdrawal date, paid on that month's settlement date)."""
        def cost(p):
            per_unit, day = costAndCashFlowDate(p)
            if per_unit < 0:
                raise ArgumentException("Per unit withdrawal cost must be non-negative.")
            return float(per_unit), day
        self._withdraw_cost = cost
        return self

    def WithWithdrawalCost(self, withdrawalCost):
        raise UnsupportedStorageFeature("WithWithdrawalCost(Func) has no device-evaluable form (CmdtyStorage.cs:411); "
                                        "use a per-unit cost.")

    # -- commodity consumed (CmdtyStorage.cs:450-509)
    def WithNoCmdtyConsumedOnInject(self):
        self._inject_consumed = lambda p: 0.0
        return self

    def WithFixedPercentCmdtyConsumedOnInject(self, percentCmdtyConsumed):
        self._inject_consumed = lambda p: float(percentCmdtyConsumed)
        return self

    def WithPercentCmdtyConsumedOnInjectTimeSeries(self, series):
        self._inject_consumed = _series_lookup(series, "Percentage of cmdty consumed on inject", self._start, self._end.Offset(-1))
        return self

    def WithCmdtyConsumedOnInject(self, fn):
        raise UnsupportedStorageFeature("WithCmdtyConsumedOnInject(Func) has no device-evaluable form (CmdtyStorage.cs:473).")

    def WithNoCmdtyConsumedOnWithdraw(self):
        self._withdraw_consumed = lambda p: 0.0
        return self

    def WithFixedPercentCmdtyConsumedOnWithdraw(self, percentCmdtyConsumed):
        self._withdraw_consumed = lambda p: float(percentCmdtyConsumed)
        return self

    def WithPercentCmdtyConsumedOnWithdrawTimeSeries(self, series):
        self._withdraw_consumed = _series_lookup(series, "Percentage of cmdty consumed on withdraw", self._start, self._end.Offset(-1))
        return self

    def WithCmdtyConsumedOnWithdraw(self, fn):
        raise UnsupportedStorageFeature("WithCmdtyConsumedOnWithdraw(Func) has no device-evaluable form (CmdtyStorage.cs:504).")

    # -- inventory loss / cost (CmdtyStorage.cs:511-568)
    def WithNoCmdtyInventoryLoss(self):
        self._inventory_loss = lambda p: 0.0
        return self

    def WithFixedPercentCmdtyInventoryLoss(self, percentCmdtyInventoryLoss):
        self._inventory_loss = lambda p: float(percentCmdtyInventoryLoss)
        return self

    def WithCmdtyInventoryLossTimeSeries(self, series):
        self._inventory_loss = _series_lookup(series, "Cmdty inventory loss", self._start, self._end.Offset(-1))
        return self

    def WithCmdtyInventoryLoss(self, fn):  # Func<T,double>: structured (depends on the period only)
        self._inventory_loss = lambda p: float(fn(p))
        return self

    def WithNoInventoryCost(self):
        self._inventory_cost = lambda p: 0.0
        return self

    def WithFixedPerUnitInventoryCost(self, perUnitCost):
        self._inventory_cost = lambda p: float(perUnitCost)
        return self

    def WithPerUnitInventoryCostTimeSeries(self, series):
        self._inventory_cost = _series_lookup(series, "Per unit inventory cost", self._start, self._end.Offset(-1))
        return self

    def WithInventoryCost(self, fn):
        raise UnsupportedStorageFeature("WithInventoryCost(Func) has no device-evaluable form (CmdtyStorage.cs:539).")

    # -- terminal state (CmdtyStorage.cs:418-428)
    def WithTerminalInventoryNpv(self, c0=0.0, c_price=0.0, c_inventory=0.0, c_price_inventory=0.0):
        """Terminal NPV = c_price_inventory*price*inventory + c_price*price + c_inventory*inventory + c0.
        The reference accepts any Func<price, inventory, npv> (CmdtyStorage.cs:418); only this bilinear family is
        device-evaluable."""
        if callable(c0):
            raise UnsupportedStorageFeature("WithTerminalInventoryNpv(Func) has no device-evaluable form (CmdtyStorage.cs:418); "
                                            "pass the bilinear coefficients.")
        self._terminal = (float(c0), float(c_price), float(c_inventory), float(c_price_inventory))
        return self

    def MustBeEmptyAtEnd(self):
        self._must_be_empty = True
        return self

    def Build(self):
        for name in ("_start", "_constraint", "_min_inventory", "_max_inventory", "_inject_cost", "_withdraw_cost",
                     "_inject_consumed", "_withdraw_consumed", "_inventory_loss", "_inventory_cost"):
            if getattr(self, name) is None:
                raise ArgumentException(f"Storage builder: {name.lstrip('_')} has not been set.")
        return CmdtyStorage(self)


class _BuilderFactory:
    def __get__(self, obj, objtype=None):
        return _StorageBuilder()

    def __call__(self):
        return _StorageBuilder()


CmdtyStorage.Builder = _BuilderFactory()
