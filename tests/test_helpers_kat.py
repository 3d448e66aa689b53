"""Known-answer tests ported from the reference's own unit tests of the sim-independent helpers on the LSMC path
(SURVEY.md §8c item 1).  Each test names the reference test it ports (files under
/root/reference/tests/Cmdty.Storage.Test/).  They run against BOTH the CPU oracle and the product library's host code
(no GPU needed): these vectors are what pins the oracle.
"""
import math

import pytest

from storage_b200 import (ArgumentException, CmdtyStorage, Day, PiecewiseLinearInjectWithdrawConstraint,
                          PolynomialInjectWithdrawConstraint, StepInjectWithdrawConstraint,
                          InventoryConstraintsCannotBeFulfilledException)
from tests import kat_impl

TOL = 1e-10  # StorageHelperTest.cs:36 NumericalTolerance


@pytest.fixture(params=["oracle", "lib"])
def impl(request):
    return kat_impl.make(request.param)


# ---------------------------------------------------------------- StorageHelperTest.cs:37-259 bang-bang decision set
UNC = (-15.5, 65.685)


def test_decision_set_unconstrained_no_extra(impl):  # StorageHelperTest.cs:39-55
    assert impl.decision_set(*UNC, 1010.0, 10.0, 900.0, 1070.0, TOL, 0) == [-15.5, 0.0, 65.685]


def test_decision_set_unconstrained_extra(impl):  # :57-75
    assert impl.decision_set(*UNC, 1010.0, 10.0, 900.0, 1070.0, TOL, 1) == [-15.5, -15.5 / 2.0, 0.0, 65.685 / 2.0, 65.685]


def test_decision_set_both_positive(impl):  # :77-93
    assert impl.decision_set(15.5, 65.685, 1010.0, 10.0, 900.0, 1070.0, TOL, 0) == [15.5, 65.685]


def test_decision_set_both_positive_extra(impl):  # :95-112
    assert impl.decision_set(15.5, 65.685, 1010.0, 10.0, 900.0, 1070.0, TOL, 1) == [15.5, (15.5 + 65.685) / 2.0, 65.685]


def test_decision_set_both_negative(impl):  # :114-130
    assert impl.decision_set(-65.685, -41.5, 1000.0, 10.0, 900.0, 950.0, TOL, 0) == [-65.685, -41.5]


def test_decision_set_both_negative_extra(impl):  # :132-150
    assert impl.decision_set(-65.685, -41.5, 1000.0, 10.0, 900.0, 950.0, TOL, 1) == [-65.685, (-65.685 + -41.5) / 2.0, -41.5]


def test_decision_set_constrained_around_current(impl):  # :152-172
    cur, loss, nmin, nmax = 1010.0, 10.0, 991.87, 1051.8
    w = nmax - cur + loss
    i = nmin - cur + loss
    assert impl.decision_set(*UNC, cur, loss, nmin, nmax, TOL, 0) == [i, 0.0, w]


def test_decision_set_constrained_around_current_extra(impl):  # :174-194
    cur, loss, nmin, nmax = 1010.0, 10.0, 991.87, 1051.8
    w = nmax - cur + loss
    i = nmin - cur + loss
    assert impl.decision_set(*UNC, cur, loss, nmin, nmax, TOL, 1) == [i, i / 2.0, 0.0, w / 2.0, w]


def test_decision_set_injection_below_current(impl):  # :196-214
    cur, loss, nmin, nmax = 1010.0, 10.0, 900.0, 995.8
    assert impl.decision_set(*UNC, cur, loss, nmin, nmax, TOL, 0) == [-15.5, nmax - cur + loss]


def test_decision_set_injection_below_current_extra(impl):  # :216-236
    cur, loss, nmin, nmax = 1010.0, 10.0, 900.0, 995.8
    inj = nmax - cur + loss
    assert impl.decision_set(*UNC, cur, loss, nmin, nmax, TOL, 1) == [-15.5, (-15.5 + inj) / 2.0, inj]


def test_decision_set_withdrawal_above_current(impl):  # :238-256
    cur, loss, nmin, nmax = 1010.0, 10.0, 1001.8, 1009.51
    assert impl.decision_set(*UNC, cur, loss, nmin, nmax, TOL, 0) == [nmin - cur + loss, nmax - cur + loss]


def test_decision_set_withdrawal_above_current_extra(impl):  # :258-259 + following
    cur, loss, nmin, nmax = 1010.0, 10.0, 1001.8, 1009.51
    w, i = nmax - cur + loss, nmin - cur + loss
    assert impl.decision_set(*UNC, cur, loss, nmin, nmax, TOL, 1) == [i, (w + i) / 2.0, w]


def test_decision_set_argument_errors(impl):  # StorageHelper.cs:112-115,131,155
    with pytest.raises(impl.ArgumentError):
        impl.decision_set(*UNC, 1010.0, 10.0, 1000.0, 900.0, TOL, 0)
    with pytest.raises(impl.ArgumentError):
        impl.decision_set(*UNC, 1010.0, 10.0, 900.0, 1070.0, TOL, -1)
    with pytest.raises(impl.ArgumentError):  # max withdrawal still above next max by more than the tolerance
        impl.decision_set(-1.0, 5.0, 1010.0, 0.0, 900.0, 1000.0, TOL, 0)
    with pytest.raises(impl.ArgumentError):  # max injection still below next min
        impl.decision_set(-1.0, 5.0, 100.0, 0.0, 900.0, 1000.0, TOL, 0)


def test_decision_set_tolerance_branches(impl):  # StorageHelper.cs:124-128,148-152 ("TODO unit test code reaching here")
    # breached by less than the tolerance: snapped onto the next-step bound
    d = impl.decision_set(-1.0, 5.0, 1001.0 + 5e-11, 0.0, 900.0, 1000.0, TOL, 0)
    assert d[0] == 1000.0 - (1001.0 + 5e-11)
    d = impl.decision_set(-1.0, 5.0, 895.0 - 5e-11, 0.0, 900.0, 1000.0, TOL, 0)
    assert d[-1] == 900.0 - (895.0 - 5e-11)


# ---------------------------------------------------------------- StorageHelperTest.cs:262-278 MaxValueAndIndex
def test_max_value_and_index():
    from oracle import oracle
    assert oracle.max_value_and_index([4.5, -1.2, 6.8, 3.2]) == (6.8, 2)
    assert oracle.max_value_and_index([1.0, 7.0, 7.0, 3.0]) == (7.0, 1)  # first maximum wins (strict >, StorageHelper.cs:214)


# ---------------------------------------------------------------- StorageHelperTest.cs:280-427 CalculateInventorySpace
def _simple_storage(start, end, loss, terminal_zero=False):
    b = (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
         .WithConstantInjectWithdrawRange(-6.0, 5.0)
         .WithConstantMinInventory(0.0)
         .WithConstantMaxInventory(23.5)
         .WithPerUnitInjectionCost(1.5)
         .WithNoCmdtyConsumedOnInject()
         .WithPerUnitWithdrawalCost(0.8)
         .WithNoCmdtyConsumedOnWithdraw()
         .WithFixedPercentCmdtyInventoryLoss(loss)
         .WithNoInventoryCost())
    b = b.WithTerminalInventoryNpv(0.0) if terminal_zero else b.MustBeEmptyAtEnd()
    return b.Build()


def test_inventory_space_current_after_start(impl):  # StorageHelperTest.cs:282-349
    inj, wd, inv0, loss, mn, mx = 5.0, 6.0, 8.0, 0.03, 0.0, 23.5
    start, end, cur = Day(2019, 8, 1), Day(2019, 8, 28), Day(2019, 8, 20)
    st = _simple_storage(start, end, loss, terminal_zero=True)
    flat = st.flatten(max(start, cur), lambda d: 1.0)
    lo, hi = impl.inventory_space(flat.struct, inv0)
    assert len(lo) == end.OffsetFrom(cur)
    e_lo = inv0 * (1 - loss) - wd
    e_hi = inv0 * (1 - loss) + inj
    # the reference test recomputes the recursion as x*(1-loss)+rate where the code does x - loss*x + rate; xUnit
    # compares doubles exactly there, here one ulp of slack is allowed for the re-association
    assert lo[0] == pytest.approx(e_lo, abs=1e-13) and hi[0] == pytest.approx(e_hi, abs=1e-13)
    for i in range(1, 8):
        e_lo = max(e_lo * (1 - loss) - wd, mn)
        e_hi = min(e_hi * (1 - loss) + inj, mx)
        assert lo[i] == pytest.approx(e_lo, abs=1e-12) and hi[i] == pytest.approx(e_hi, abs=1e-12)


def test_inventory_space_current_before_start(impl):  # StorageHelperTest.cs:351-420
    inj, wd, inv0, loss, mn, mx = 5.0, 6.0, 11.0, 0.03, 0.0, 23.5
    start, end, cur = Day(2019, 8, 19), Day(2019, 8, 28), Day(2019, 8, 10)
    st = _simple_storage(start, end, loss)
    flat = st.flatten(max(start, cur), lambda d: 1.0)
    lo, hi = impl.inventory_space(flat.struct, inv0)
    assert len(lo) == end.OffsetFrom(start)
    e_lo = inv0 * (1 - loss) - wd
    e_hi = inv0 * (1 - loss) + inj
    assert lo[0] == pytest.approx(e_lo, abs=1e-13) and hi[0] == pytest.approx(e_hi, abs=1e-13)
    for i in range(1, 5):  # 21st .. 24th
        e_lo = max(e_lo * (1 - loss) - wd, mn)
        e_hi = min(e_hi * (1 - loss) + inj, mx)
        assert lo[i] == pytest.approx(e_lo, abs=1e-12) and hi[i] == pytest.approx(e_hi, abs=1e-12)
    # backward-derived part: 28th is (0,0), then upper = min((upper + wd)/(1-loss), max)
    assert (lo[8], hi[8]) == (0.0, 0.0)
    e_hi = 0.0
    for i in (7, 6, 5):
        e_hi = min((e_hi + wd) / (1 - loss), mx)
        assert lo[i] == 0.0 and hi[i] == pytest.approx(e_hi, abs=1e-12)


def test_inventory_space_cannot_be_fulfilled(impl):  # StorageHelper.cs:101-102
    start, end = Day(2019, 8, 19), Day(2019, 8, 22)
    st = _simple_storage(start, end, 0.0)  # must be empty after 3 days, can only withdraw 6/day
    flat = st.flatten(start, lambda d: 1.0)
    with pytest.raises((InventoryConstraintsCannotBeFulfilledException, Exception)) as e:
        impl.inventory_space(flat.struct, 23.0)
    assert getattr(e.value, "code", 2) == 2 or isinstance(e.value, InventoryConstraintsCannotBeFulfilledException)


# ---------------------------------------------------------------- StorageHelperTest.cs:429-607 BisectInventorySpace
GRID5 = [0.0, 5.3, 9.5, 15.63, 25.8]


@pytest.mark.parametrize("grid,inv,expected", [
    (GRID5, 25.8, (3, 4)),               # InventoryEqualsHighestGridValue_ReturnsTopIndexMinusOneAndTopIndex
    (GRID5, 0.0, (0, 0)),                # InventoryEqualsLowestGridValue_ReturnsZeroAndZero
    (GRID5, 20.01, (3, 4)),              # InventoryBetweenTopTwoValues
    (GRID5, 1.23, (0, 1)),               # InventoryBetweenBottomTwoValues
    (GRID5, 5.3, (1, 1)),                # InventoryEqualsSecondLowestValue_ReturnsOneAndOne
    (GRID5, 10.89, (2, 3)),              # InventoryWithinGrid_AsExpected
    ([1.3, 2.8], 1.4, (0, 1)),           # GridWithTwoPointsInventoryWithin
    ([1.3, 2.8], 1.3, (0, 0)),           # GridWithTwoPointsInventoryEqualsLowestPoint
    ([1.3, 2.8], 2.8, (0, 1)),           # GridWithTwoPointsInventoryEqualsUpperPoint_ReturnsZeroAndOne
    ([1.3], 1.3, (0, 0)),                # SingleItemGridEqualsInventory
    (GRID5, 0.0 - TOL / 2.0, (0, 0)),    # InventoryLowerButWithinToleranceOfInventoryGridMin
    (GRID5, 25.8 + TOL / 2.0, (4, 4)),   # InventoryHigherButWithinToleranceOfInventoryGridMax
])
def test_bisect_inventory_space(impl, grid, inv, expected):
    assert impl.bisect(grid, inv, TOL) == expected


@pytest.mark.parametrize("inv", [26.8, -1.2])  # InventoryHigherThanRange_Throws / InventoryLowerThanRange_Throws
def test_bisect_inventory_space_throws(impl, inv):
    with pytest.raises(impl.ArgumentError):
        impl.bisect(GRID5, inv, TOL)


def test_bisect_matches_linear_definition(impl):
    """Property: on a long ragged grid the bisection brackets every interior point and returns exact hits."""
    import numpy as np
    rng = np.random.default_rng(5)
    grid = np.cumsum(rng.uniform(0.5, 3.0, 257))
    for x in rng.uniform(grid[0], grid[-1], 500):
        lo, hi = impl.bisect(grid, float(x), TOL)
        assert hi == lo + 1 and grid[lo] <= x <= grid[hi]
    for i in (0, 1, 100, 255, 256):
        lo, hi = impl.bisect(grid, float(grid[i]), TOL)
        # an exact hit is reported as (i,i) unless the search meets it first as the upper end of a bracket
        assert (lo, hi) == (i, i) or (lo, hi) == (i - 1, i)


# ---------------------------------------------------------------- FixedSpacingStateSpaceGridCalcTest.cs:35-135
def test_grid_space_less_than_spacing(impl):
    assert impl.grid_points(120.55, 161.89, 50.0) == [120.55, 161.89]


def test_grid_upper_equals_lower_plus_spacing(impl):
    assert impl.grid_points(121.2, 121.2 + 15.0, 15.0) == [121.2, 121.2 + 15.0]


def test_grid_separated_by_spacing(impl):  # :70-89
    assert impl.grid_points(121.2, 174.89, 15.0) == [121.2, 136.2, 151.2, 166.2, 174.89]


def test_grid_single_point(impl):
    assert impl.grid_points(121.2, 121.2, 15.0) == [121.2]


def test_grid_lower_above_upper_throws(impl):
    with pytest.raises(impl.ArgumentError):
        impl.grid_points(121.2, 121.19, 15.0)


def test_grid_calc_constructor_validation():
    from storage_b200 import FixedSpacingStateSpaceGridCalc
    with pytest.raises(ArgumentException):
        FixedSpacingStateSpaceGridCalc(-15.0)
    with pytest.raises(ArgumentException):
        FixedSpacingStateSpaceGridCalc(0.0)


def test_grid_accumulates_additions(impl):
    """Grid points are lower + spacing + spacing + ... (accumulated, FixedSpacingStateSpaceGridCalc.cs:54-60), not lower + i*spacing."""
    spacing = 52500.0 / 99
    pts = impl.grid_points(0.0, 52500.0, spacing)
    acc, expected = 0.0, [0.0]
    while acc < 52500.0:
        acc += spacing
        expected.append(min(acc, 52500.0))
    assert pts == expected


# ---------------------------------------------------------------- constraint classes
RATCHETS = [(0.0, (-44.85, 56.8)), (100.0, (-45.01, 54.5)), (300.0, (-45.78, 52.01)), (600.0, (-46.17, 51.9)),
            (800.0, (-46.99, 50.8)), (1000.0, (-47.12, 50.01))]
CONST_ROWS = [(x, (-47.12, 56.8)) for x in (0.0, 100.0, 300.0, 600.0, 800.0, 1000.0)]


def _storage_with(constraint, loss, mn=0.0, mx=1000.0):
    start = Day(2020, 1, 1)
    st = (CmdtyStorage.Builder.WithActiveTimePeriod(start, start + 3)
          .WithInjectWithdrawConstraint(constraint)
          .WithConstantMinInventory(mn).WithConstantMaxInventory(mx)
          .WithPerUnitInjectionCost(0.0).WithNoCmdtyConsumedOnInject()
          .WithPerUnitWithdrawalCost(0.0).WithNoCmdtyConsumedOnWithdraw()
          .WithFixedPercentCmdtyInventoryLoss(loss).WithNoInventoryCost()
          .WithTerminalInventoryNpv(0.0).Build())
    return st.flatten(start, lambda d: 1.0)


def _const(min_rate, max_rate):
    from storage_b200 import ConstantInjectWithdrawConstraint
    return ConstantInjectWithdrawConstraint(min_rate, max_rate)


def test_constant_constraint_bounds_consistent(impl):  # ConstantInjectWithdrawConstraintTest.cs:33-79
    loss = 0.03
    flat = _storage_with(_const(-44.85, 56.8), loss)
    lo, hi = impl.space_bounds(flat.struct, 0, 320.0, 620.0)
    assert hi * (1 - loss) + impl.range(flat.struct, 0, hi)[0] == pytest.approx(620.0, abs=1e-12)
    assert lo * (1 - loss) + impl.range(flat.struct, 0, lo)[1] == pytest.approx(320.0, abs=1e-12)


@pytest.mark.parametrize("factory", [PiecewiseLinearInjectWithdrawConstraint, StepInjectWithdrawConstraint,
                                     PolynomialInjectWithdrawConstraint])
def test_constraint_constant_rates_upper(impl, factory):  # PiecewiseLinear…Test.cs:32-57, Step…Test.cs:50-75, Polynomial…Test.cs:32-57
    loss = 0.03
    flat = _storage_with(factory(CONST_ROWS), loss)
    _, hi = impl.space_bounds(flat.struct, 0, 320.0, 620.0)
    assert hi == pytest.approx((620.0 + 47.12) / (1 - loss), abs=1e-10 if factory is PolynomialInjectWithdrawConstraint else 1e-12)


@pytest.mark.parametrize("factory", [PiecewiseLinearInjectWithdrawConstraint, StepInjectWithdrawConstraint,
                                     PolynomialInjectWithdrawConstraint])
def test_constraint_constant_rates_lower(impl, factory):  # PiecewiseLinear…Test.cs:59-84 and the Step/Polynomial twins
    loss = 0.03
    flat = _storage_with(factory(CONST_ROWS), loss)
    lo, _ = impl.space_bounds(flat.struct, 0, 620.0, 870.0)
    assert lo == pytest.approx((620.0 - 56.8) / (1 - loss), abs=1e-10 if factory is PolynomialInjectWithdrawConstraint else 1e-12)


def test_piecewise_linear_bounds_consistent(impl):  # PiecewiseLinear…Test.cs:86-141
    loss = 0.03
    flat = _storage_with(PiecewiseLinearInjectWithdrawConstraint(RATCHETS), loss)
    _, hi = impl.space_bounds(flat.struct, 0, 320.0, 620.0)
    assert hi * (1 - loss) + impl.range(flat.struct, 0, hi)[0] == pytest.approx(620.0, abs=1e-12)
    lo, _ = impl.space_bounds(flat.struct, 0, 552.0, 734.0)
    assert lo * (1 - loss) + impl.range(flat.struct, 0, lo)[1] == pytest.approx(552.0, abs=1e-12)


def test_piecewise_linear_step_function_approximation(impl):  # PiecewiseLinear…Test.cs:143-200
    rows = [(0.0, (-5810.0, 7632.0)), (64000.0, (-5810.0, 7632.0)), (64260.0, (-15288.0, 7632.0)),
            (182070.0, (-15288.0, 3816.0)), (192780.0, (-15288.0, 2279.0)), (214200.0, (-15288.0, 2269.0))]
    loss = 0.03
    flat = _storage_with(PiecewiseLinearInjectWithdrawConstraint(rows), loss, 0.0, 214200.0)
    _, hi = impl.space_bounds(flat.struct, 0, 320.0, 65260.0)
    assert hi * (1 - loss) + impl.range(flat.struct, 0, hi)[0] == pytest.approx(65260.0, abs=1e-9)
    rows[1] = (64259.0, (-5810.0, 7632.0))
    flat = _storage_with(PiecewiseLinearInjectWithdrawConstraint(rows), 0.0, 0.0, 214200.0)
    lo, _ = impl.space_bounds(flat.struct, 0, 65260.0, 172070.0)
    assert lo + impl.range(flat.struct, 0, lo)[1] == pytest.approx(65260.0, abs=1e-9)


def test_piecewise_linear_range_at_pillars_exact(impl):  # PiecewiseLinear…Test.cs:202-223 (exact equality)
    flat = _storage_with(PiecewiseLinearInjectWithdrawConstraint(RATCHETS), 0.0)
    for inv, (mn, mx) in RATCHETS:
        assert impl.range(flat.struct, 0, inv) == (mn, mx)


def test_piecewise_linear_range_halfway(impl):  # PiecewiseLinear…Test.cs:225-245 (exact equality)
    flat = _storage_with(PiecewiseLinearInjectWithdrawConstraint(RATCHETS), 0.0)
    assert impl.range(flat.struct, 0, 200.0) == ((-45.01 - 45.78) / 2.0, (54.5 + 52.01) / 2.0)


STEP_ROWS = RATCHETS[:5] + [(1000.0, (-46.99, 50.8))]


def test_step_bounds_consistent(impl):  # Step…Test.cs:104-131
    loss = 0.03
    flat = _storage_with(StepInjectWithdrawConstraint(STEP_ROWS), loss)
    _, hi = impl.space_bounds(flat.struct, 0, 320.0, 590.5)
    assert hi * (1 - loss) + impl.range(flat.struct, 0, hi)[0] == pytest.approx(590.5, abs=1e-12)
    lo, _ = impl.space_bounds(flat.struct, 0, 552.0, 734.0)
    assert lo * (1 - loss) + impl.range(flat.struct, 0, lo)[1] == pytest.approx(552.0, abs=1e-12)


@pytest.mark.parametrize("inv,mn,mx", [(99.0, -44.85, 56.8), (610.85, -46.17, 51.9), (999.99, -46.99, 50.8)])
def test_step_range_as_expected(impl, inv, mn, mx):  # Step…Test.cs:133-143 [InlineData]
    flat = _storage_with(StepInjectWithdrawConstraint(STEP_ROWS), 0.0)
    assert impl.range(flat.struct, 0, inv) == (mn, mx)


def test_step_range_at_pillars(impl):  # Step…Test.cs:145-154
    flat = _storage_with(StepInjectWithdrawConstraint(STEP_ROWS), 0.0)
    for inv, (mn, mx) in STEP_ROWS:
        assert impl.range(flat.struct, 0, inv) == (mn, mx)


def test_step_range_outside_throws(impl):  # Step…Test.cs:156-166 (CmdtyStorage.cs:88-94 throws first through the storage)
    flat = _storage_with(StepInjectWithdrawConstraint(STEP_ROWS), 0.0)
    with pytest.raises(impl.ArgumentError):
        impl.range(flat.struct, 0, -0.01)
    with pytest.raises(impl.ArgumentError):
        impl.range(flat.struct, 0, 1000.001)


def test_step_constructor_validation():  # Step…Test.cs:168-198
    bad = list(STEP_ROWS)
    bad[3] = (600.0, (-46.17, 52.01001))
    with pytest.raises(ArgumentException):
        StepInjectWithdrawConstraint(bad)
    bad = list(STEP_ROWS)
    bad[3] = (600.0, (-45.77, 51.9))
    with pytest.raises(ArgumentException):
        StepInjectWithdrawConstraint(bad)


def test_polynomial_bounds_consistent(impl):  # Polynomial…Test.cs:86-141 (root finder accuracy 1e-10)
    loss = 0.03
    flat = _storage_with(PolynomialInjectWithdrawConstraint(RATCHETS), loss)
    _, hi = impl.space_bounds(flat.struct, 0, 320.0, 620.0)
    assert hi * (1 - loss) + impl.range(flat.struct, 0, hi)[0] == pytest.approx(620.0, abs=1e-9)
    lo, _ = impl.space_bounds(flat.struct, 0, 552.0, 734.0)
    assert lo * (1 - loss) + impl.range(flat.struct, 0, lo)[1] == pytest.approx(552.0, abs=1e-9)


def test_polynomial_interpolates_pillars(impl):
    flat = _storage_with(PolynomialInjectWithdrawConstraint(RATCHETS), 0.0)
    for inv, (mn, mx) in RATCHETS:
        r = impl.range(flat.struct, 0, inv)
        assert r[0] == pytest.approx(mn, abs=1e-8) and r[1] == pytest.approx(mx, abs=1e-8)


def test_storage_range_outside_inventory_limits_throws(impl):  # CmdtyStorage.cs:88-94
    flat = _storage_with(_const(-1.0, 1.0), 0.0, 10.0, 20.0)
    with pytest.raises(impl.ArgumentError):
        impl.range(flat.struct, 0, 9.99)
    with pytest.raises(impl.ArgumentError):
        impl.range(flat.struct, 0, 20.01)
    assert impl.range(flat.struct, 3, 15.0) == (0.0, 0.0)  # period >= EndPeriod: (0,0) (CmdtyStorage.cs:96-97)


# ---------------------------------------------------------------- CmdtyStorageBuilderTest.cs / builder validation
def test_builder_validation():
    with pytest.raises(ArgumentException):
        CmdtyStorage.Builder.WithActiveTimePeriod(Day(2020, 1, 2), Day(2020, 1, 2))
    b = CmdtyStorage.Builder.WithActiveTimePeriod(Day(2020, 1, 1), Day(2020, 1, 5)).WithConstantInjectWithdrawRange(-1, 1)
    with pytest.raises(ArgumentException):
        b.WithConstantMinInventory(-1.0)
    with pytest.raises(ArgumentException):
        b.WithConstantMaxInventory(-1.0)
    with pytest.raises(ArgumentException):
        b.WithPerUnitInjectionCost(-0.1)
    with pytest.raises(ArgumentException):
        b.WithPerUnitWithdrawalCost(-0.1)


def test_ratchets_change_in_time_and_two_equal_rows_are_constant():  # CmdtyStorageBuilderExtensions.cs:142-257
    from storage_b200 import _abi
    start = Day(2019, 9, 1)
    st = (CmdtyStorage.Builder.WithActiveTimePeriod(start, Day(2019, 10, 1))
          .WithTimeAndInventoryVaryingInjectWithdrawRatesPiecewiseLinear([
              (start, RATCHETS),
              (Day(2019, 9, 10), [(0.0, (0.0, 0.0)), (1000.0, (0.0, 0.0))]),  # outage: constant (0,0)
              (Day(2019, 9, 20), RATCHETS)])
          .WithPerUnitInjectionCost(0.48).WithNoCmdtyConsumedOnInject()
          .WithPerUnitWithdrawalCost(0.74).WithNoCmdtyConsumedOnWithdraw()
          .WithNoCmdtyInventoryLoss().WithNoInventoryCost().MustBeEmptyAtEnd().Build())
    flat = st.flatten(start, lambda d: 1.0)
    kinds = list(flat.arrays["constraint_kind"])
    assert kinds[0] == _abi.LSMC_CONSTRAINT_PIECEWISE_LINEAR
    assert kinds[9] == _abi.LSMC_CONSTRAINT_CONSTANT and kinds[18] == _abi.LSMC_CONSTRAINT_CONSTANT
    assert kinds[19] == _abi.LSMC_CONSTRAINT_PIECEWISE_LINEAR
    assert st.MinInventory(Day(2019, 9, 15)) == 0.0 and st.MaxInventory(Day(2019, 9, 15)) == 1000.0
    assert st.MaxInventory(Day(2019, 10, 1)) == 0.0  # MustBeEmptyAtEnd (CmdtyStorage.cs:435-438)


def test_act365_discounter():  # StorageHelper.cs:262-273
    from storage_b200 import StorageHelper
    d = StorageHelper.CreateAct65ContCompDiscounter(0.055)
    assert d(Day(2019, 8, 29), Day(2019, 8, 29)) == 1.0
    assert d(Day(2019, 8, 29), Day(2019, 8, 1)) == 1.0
    assert d(Day(2019, 8, 29), Day(2019, 9, 20)) == math.exp(-22 / 365.0 * 0.055)
