"""The oracle's shared machinery pinned on the reference's own golden values.

IntrinsicStorageValuation (IntrinsicValuation/IntrinsicStorageValuation.cs:120-322) is deterministic and uses the same
CalculateInventorySpace / grid / CalculateBangBangDecisionSet / MaxValueAndIndex helpers, storage entity, settlement and
discounting conventions as the LSMC valuation.  The reference's Python regression test holds exact goldens for it:
`intrinsic_npv == 1703773.0757192627` and `regression_test_data/multi_factor_test-1/intrinsic_profile.csv`
(src/Cmdty.Storage.Python/tests/test_multi_factor.py:34-62,102,123).  The inputs below are that test's inputs; the CSV is the
reference's fixture, copied verbatim to tests/golden/.
"""
import csv
import os

import numpy as np
import pytest

from oracle import oracle
from storage_b200 import (BasisFunctionsBuilder, CmdtyStorage, Day, FixedSpacingStateSpaceGridCalc, FlatProblem,
                          LsmcValuationParameters, MultiFactorParameters, StorageHelper, settle_on_20th_of_following_month)

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "reference_multi_factor_test-1_intrinsic_profile.csv")
REFERENCE_INTRINSIC_NPV = 1703773.0757192627  # test_multi_factor.py:102 (assertEqual: exact)


def reference_problem():
    start, end = Day(2019, 12, 1), Day(2020, 4, 1)
    storage = (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
               .WithConstantInjectWithdrawRange(-700.0, 700.0)  # cmdty_storage.py:116
               .WithConstantMinInventory(0.0)
               .WithConstantMaxInventory(100000.0)
               .WithPerUnitInjectionCost(1.23)
               .WithNoCmdtyConsumedOnInject()
               .WithPerUnitWithdrawalCost(0.98)
               .WithNoCmdtyConsumedOnWithdraw()
               .WithNoCmdtyInventoryLoss()
               .WithNoInventoryCost()
               .MustBeEmptyAtEnd()
               .Build())
    val_date = Day(2019, 8, 29)
    switch = Day(2020, 3, 12)
    fwd = {d: (23.87 if d < switch else 150.32) for d in val_date.EnumerateTo(end)}
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = 0.0
    b.ForwardCurve = fwd
    b.Storage = storage
    b.SettleDateRule = settle_on_20th_of_following_month
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.03)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 100)
    b.NumericalTolerance = 1e-12  # intrinsic.py:40
    b.BasisFunctions = BasisFunctionsBuilder.Ones  # not used by the intrinsic valuation
    b.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(1.0, 0.1), 16, 1)
    return FlatProblem(b.Build(), 16, 1)


def load_golden():
    with open(GOLDEN, newline="") as f:
        rows = list(csv.reader(f))
    assert rows[0][1:] == ["inventory", "inject_withdraw_volume", "cmdty_consumed", "inventory_loss", "net_volume", "period_pv"]
    dates = [r[0] for r in rows[1:]]
    data = np.array([[float(x) for x in r[1:]] for r in rows[1:]])
    return dates, data


def test_intrinsic_npv_equals_the_reference_golden():
    npv, _ = oracle.intrinsic_calculate(reference_problem())
    assert npv == REFERENCE_INTRINSIC_NPV  # bit for bit


def test_intrinsic_profile_equals_the_reference_golden():
    flat = reference_problem()
    _, profile = oracle.intrinsic_calculate(flat)
    dates, golden = load_golden()
    assert len(dates) == flat.num_periods + 1 == 123
    assert dates[0] == "2019-12-01" and dates[-1] == "2020-04-01"
    np.testing.assert_array_equal(profile[:, 0], golden[:, 0])  # inventory
    np.testing.assert_array_equal(profile[:, 1], golden[:, 1])  # inject / withdraw volume
    np.testing.assert_array_equal(profile[:, 2], golden[:, 2])  # commodity consumed
    np.testing.assert_array_equal(profile[:, 3], golden[:, 3])  # inventory loss
    np.testing.assert_array_equal(-profile[:, 1] - profile[:, 2], golden[:, 4])  # net volume (StorageProfile.NetVolume)
    np.testing.assert_array_equal(profile[:, 4], golden[:, 5])  # period PV, bit for bit


def readme_problem():
    """README.md:322-350,400-440 = samples/csharp/Cmdty.Storage.Samples.Intrinsic/Program.cs:31-81: constant-parameter storage,
    valuation inside the active period, no discounting, grid spacing 10; the program prints the NPV as 10,827.21."""
    storage = (CmdtyStorage.Builder.WithActiveTimePeriod(Day(2019, 9, 1), Day(2019, 10, 1))
               .WithConstantInjectWithdrawRange(-14.74, 5.26)
               .WithConstantMinInventory(0.0)
               .WithConstantMaxInventory(1100.74)
               .WithPerUnitInjectionCost(0.48, lambda d: d)
               .WithNoCmdtyConsumedOnInject()
               .WithPerUnitWithdrawalCost(0.74, lambda d: d)
               .WithNoCmdtyConsumedOnWithdraw()
               .WithNoCmdtyInventoryLoss()
               .WithNoInventoryCost()
               .MustBeEmptyAtEnd()
               .Build())
    current = Day(2019, 9, 15)
    fwd = {d: 56.6 for d in Day(2019, 9, 15).EnumerateTo(Day(2019, 9, 22))}
    fwd.update({d: 56.6 + 87.81 for d in Day(2019, 9, 23).EnumerateTo(Day(2019, 10, 1))})
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = current
    b.Inventory = 50.0
    b.ForwardCurve = fwd
    b.Storage = storage
    b.SettleDateRule = settle_on_20th_of_following_month  # irrelevant: no discounting
    b.DiscountFactors = lambda present, cash_flow: 1.0
    b.GridCalc = FixedSpacingStateSpaceGridCalc(10.0)
    b.NumericalTolerance = 1e-12
    b.BasisFunctions = BasisFunctionsBuilder.Ones
    b.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(1.0, 0.1), 16, 1)
    return FlatProblem(b.Build(), 16, 1)


def test_readme_intrinsic_value():
    npv, profile = oracle.intrinsic_calculate(readme_problem())
    assert f"{npv:,.2f}" == "10,827.21"  # README.md:436-440 (Npv.ToString("N2"))
    assert profile[-1, 0] == pytest.approx(0.0, abs=1e-9)  # must be empty at the end


# ---------------------------------------------------------------------------------------------------------------------
# The CUDA valuation against the golden-pinned intrinsic valuation: with a very small volatility the LSMC value must
# collapse onto the intrinsic value (LsmcStorageValuationTest.cs:573-610, 616-653; same fixture, tolerances and path count;
# the paths come from the library's own simulator).
# ---------------------------------------------------------------------------------------------------------------------
def _reference_fixture(storage_kind):
    import math
    from storage_b200 import LsmcStorageValuation
    start, end = Day(2019, 8, 3), Day(2020, 4, 1)  # LsmcStorageValuationTest.cs:71-72
    b = CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
    if storage_kind == "simple":  # :81-93
        b = b.WithConstantInjectWithdrawRange(-850.0, 625.0).WithZeroMinInventory().WithConstantMaxInventory(52_500.0)
    else:  # :95-123
        b = b.WithTimeAndInventoryVaryingInjectWithdrawRatesPiecewiseLinear([
            (start, [(0.0, (-702.7, 650.0)), (15_000.0, (-785.0, 552.5)), (30_000.0, (-790.6, 512.8)),
                     (40_000.0, (-825.6, 498.6)), (52_500.0, (-850.4, 480.0))]),
            (Day(2020, 2, 1), [(0.0, (-645.35, 650.0)), (13_000.0, (-656.0, 552.5)), (28_000.0, (-689.6, 512.8)),
                               (42_000.0, (-701.06, 498.6)), (52_500.0, (-718.04, 480.0))])])
    storage = (b.WithPerUnitInjectionCost(1.25, lambda d: d).WithNoCmdtyConsumedOnInject()
               .WithPerUnitWithdrawalCost(0.93, lambda d: d).WithNoCmdtyConsumedOnWithdraw()
               .WithNoCmdtyInventoryLoss().WithNoInventoryCost().MustBeEmptyAtEnd().Build())
    val_date = Day(2019, 8, 29)  # :126
    p = LsmcValuationParameters.Builder()
    p.CurrentPeriod = val_date
    p.Inventory = 5_685.0
    p.ForwardCurve = {d: 53.5 + math.sin(2.0 * math.pi / 365.0 * d.OffsetFrom(val_date)) * 24.6 for d in val_date.EnumerateTo(end)}
    p.Storage = storage
    p.SettleDateRule = settle_on_20th_of_following_month
    p.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.055)
    p.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 100)
    p.BasisFunctions = (BasisFunctionsBuilder.Ones + BasisFunctionsBuilder.SpotPricePower(1) + BasisFunctionsBuilder.SpotPricePower(2)
                        + BasisFunctionsBuilder.SpotPricePower(3))
    p.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(12.5, 0.001), 2_000, 11)  # SmallVol, NumSims, seed
    params = p.Build()
    lsmc = LsmcStorageValuation(device=0).Calculate(params)
    intrinsic_npv, intrinsic_profile = oracle.intrinsic_calculate(FlatProblem(params, 2_000, 1))
    return params, lsmc, intrinsic_npv, intrinsic_profile


@pytest.mark.gpu
@pytest.mark.parametrize("storage_kind,percent_tol", [("simple", 0.0001), ("ratchets", 0.0004)])
def test_very_low_vol_lsmc_npv_equals_intrinsic_npv(storage_kind, percent_tol):
    _, lsmc, intrinsic_npv, _ = _reference_fixture(storage_kind)
    assert abs(lsmc.Npv - intrinsic_npv) / abs(intrinsic_npv) <= percent_tol  # TestHelper.AssertWithinPercentTol


@pytest.mark.gpu
def test_very_low_vol_lsmc_deltas_equal_intrinsic_volume_profile():  # LsmcStorageValuationTest.cs:612-653
    params, lsmc, _, profile = _reference_fixture("simple")
    periods = sorted(lsmc.Deltas.keys())
    assert len(periods) == profile.shape[0]
    net_volume = -profile[:, 1] - profile[:, 2]
    for i, day in enumerate(periods):
        assert round(abs(lsmc.Deltas[day] - net_volume[i]), 5) == 0.0  # Assert.Equal(expected, actual, precision: 5)


# ---------------------------------------------------------------------------------------------------------------------
# The reference's closed-form intrinsic tests (tests/Cmdty.Storage.Test/IntrinsicStorageValuationTest.cs:37-183)
# ---------------------------------------------------------------------------------------------------------------------
def _intrinsic_fixture(starting_inventory, fwd, current):  # GenerateValuationResults, :37-76
    storage = (CmdtyStorage.Builder.WithActiveTimePeriod(Day(2019, 9, 1), Day(2019, 9, 30))
               .WithConstantInjectWithdrawRange(-45.5, 56.6)
               .WithConstantMinInventory(0.0)
               .WithConstantMaxInventory(1000.0)
               .WithPerUnitInjectionCost(0.8, lambda d: d)
               .WithNoCmdtyConsumedOnInject()
               .WithPerUnitWithdrawalCost(1.2, lambda d: d)
               .WithNoCmdtyConsumedOnWithdraw()
               .WithNoCmdtyInventoryLoss()
               .WithNoInventoryCost()
               .MustBeEmptyAtEnd()
               .Build())
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = current
    b.Inventory = starting_inventory
    b.ForwardCurve = fwd
    b.Storage = storage
    b.SettleDateRule = lambda d: Day(2019, 10, 5)  # WithMonthlySettlement({2019-09: 2019-10-05})
    b.DiscountFactors = lambda present, cash_flow: 1.0
    b.GridCalc = FixedSpacingStateSpaceGridCalc(10.0)
    b.NumericalTolerance = 1e-10
    b.BasisFunctions = BasisFunctionsBuilder.Ones
    b.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(1.0, 0.1), 16, 1)
    return oracle.intrinsic_calculate(FlatProblem(b.Build(), 16, 1))


def _spread_curve(spread):  # :148-171
    fwd = {d: 56.6 for d in Day(2019, 9, 15).EnumerateTo(Day(2019, 9, 22))}
    fwd.update({d: 56.6 + spread for d in Day(2019, 9, 23).EnumerateTo(Day(2019, 9, 30))})
    return fwd


def test_intrinsic_backwardated_curve_has_zero_value_and_no_flows():  # :102-128
    price, fwd = 59.95, {}
    for d in Day(2019, 9, 1).EnumerateTo(Day(2019, 9, 30)):
        fwd[d] = price
        price -= 0.58
    npv, profile = _intrinsic_fixture(0.0, fwd, Day(2019, 9, 15))
    assert npv == 0.0
    assert profile.shape[0] == 16  # 2019-09-15 .. 2019-09-30
    assert np.all(profile[:, 1] == 0.0)


def test_intrinsic_spread_below_cycle_cost_has_zero_value_and_no_flows():  # :130-146
    npv, profile = _intrinsic_fixture(0.0, _spread_curve(1.99), Day(2019, 9, 15))
    assert npv == 0.0
    assert np.all(profile[:, 1] == 0.0)


def test_intrinsic_spread_above_cycle_cost_earns_spread_minus_cost_on_the_cyclable_volume():  # :173-183
    npv, _ = _intrinsic_fixture(0.0, _spread_curve(2.01), Day(2019, 9, 15))
    assert round(abs(npv - 7 * 45.5 * 0.01), 10) == 0.0  # Assert.Equal(expected, actual, precision: 10)
