"""The reference's statistical tests of Calculate that need no particular random-number stream, run
against the CUDA library with its own device simulator: any correct simulator of the same model and any correct LSMC must
pass them.  Fixtures are the reference's (tests/Cmdty.Storage.Test/Lsmc/LsmcStorageValuationTest.cs and TestHelper.cs), the
closed forms are restated from TestHelper.cs:38-85.

  storage that is three call options          NPV vs Black-76, deltas vs Black-76 deltas      :312-423
  zero mean reversion                         NPV within 3 % of the intrinsic value             :528-571
  two factors, very low volatility            NPV within 0.001 % of the intrinsic value         :655-687
  forced injection, then forced withdrawal    NPV within 3 % of the hand calculation            :692-785

The reference runs the call-option tests on 2,000 paths of ITS stream and asserts NPV error in [-2 %, 0]; with another stream
the sampling error of a 2,000-path estimate (about 2 %) decides the sign, so they run on 100,000 paths here and the upper
bound is two standard errors instead of zero.  The intrinsic values come from the oracle's intrinsic valuation, which is
pinned bit for bit on the reference's golden value (tests/test_intrinsic_golden.py)."""
import math

import numpy as np
import pytest

from oracle import oracle
from storage_b200 import (BasisFunctionsBuilder, CmdtyStorage, Day, FixedSpacingStateSpaceGridCalc, FlatProblem, LsmcStorageValuation,
                          LsmcValuationParameters, Month, MultiFactorParameters, StorageHelper)
from tests.test_intrinsic_golden import _reference_fixture  # noqa: F401  (same storages / curve as the low-vol tests)

pytestmark = pytest.mark.gpu

ONE_FACTOR_BASIS = (BasisFunctionsBuilder.Ones + BasisFunctionsBuilder.SpotPricePower(1) + BasisFunctionsBuilder.SpotPricePower(2)
                    + BasisFunctionsBuilder.SpotPricePower(3))


def norm_cdf(x):
    return 0.5 * (1.0 + math.erf(x / math.sqrt(2.0)))


def black76_call(val_date, forward, implied_vol, rate, strike, expiry, settle):  # TestHelper.cs:38-54
    df = math.exp(-settle.OffsetFrom(val_date) / 365.0 * rate)
    t = expiry.OffsetFrom(val_date) / 365.0
    vt = implied_vol * math.sqrt(t)
    d1 = (math.log(forward / strike) + implied_vol * implied_vol / 2.0 * t) / vt
    return df * (forward * norm_cdf(d1) - strike * norm_cdf(d1 - vt)), norm_cdf(d1)


def one_factor_implied_vol(val_date, expiry, spot_vol, mean_reversion):  # TestHelper.cs:76-85
    t = expiry.OffsetFrom(val_date) / 365.0
    variance = (1.0 - math.exp(-2.0 * mean_reversion * t)) / 2.0 / mean_reversion
    return spot_vol[expiry] * math.sqrt(variance / t)


def make_test_curves(start, end):  # TestHelper.CreateDailyTestForwardAndSpotVolCurves, TestHelper.cs:87-108
    fwd, vol = {}, {}
    for i, d in enumerate(start.EnumerateTo(end)):
        fwd[d] = 53.5 + math.sin(2.0 * math.pi / 365.0 * i) * 24.6
        vol[d] = 0.78 + math.sin(2.0 * math.pi / 365.0 * i) * 0.35
    return fwd, vol


SETTLE = {Month(2019, 12): Day(2020, 1, 20), Month(2020, 1): Day(2020, 2, 18), Month(2020, 2): Day(2020, 3, 21), Month(2020, 3): Day(2020, 4, 22)}


def settle_rule(day):
    return SETTLE[Month(day.Start.year, day.Start.month)]


def three_calls(fwd):  # TestHelper.CreateThreeCallsLikeStorageTestData, TestHelper.cs:110-210
    dates = [Day(2019, 12, 15), Day(2020, 1, 20), Day(2020, 3, 31)]
    notionals = [1200.0, 800.0, 900.0]
    strikes = [fwd[dates[0]], fwd[dates[1]] + 2.0, fwd[dates[2]] + 2.8]
    start, end = Day(2019, 12, 1), Day(2020, 4, 1)
    inventory = sum(notionals)
    closed = [(0.0, (0.0, 0.0)), (inventory, (0.0, 0.0))]
    sched = [(start, closed)]
    for d, n in zip(dates, notionals):
        sched += [(d, [(0.0, (-n, 0.0)), (inventory, (-n, 0.0))]), (d + 1, closed)]
    strike_of = dict(zip(dates, strikes))
    storage = (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
               .WithTimeAndInventoryVaryingInjectWithdrawRatesPolynomial(sched)
               .WithPerUnitInjectionCost(0.0, lambda d: d).WithNoCmdtyConsumedOnInject()
               .WithPerUnitWithdrawalCostByPeriod(lambda d: (strike_of.get(d, 0.0), settle_rule(d)))  # WithWithdrawalCost(Func), TestHelper.cs:177-195
               .WithNoCmdtyConsumedOnWithdraw().WithNoCmdtyInventoryLoss().WithNoInventoryCost()
               .WithTerminalInventoryNpv(0.0).Build())  # value in storage at the end is worthless
    calls = [dict(expiry=d, strike=k, notional=n, settle=settle_rule(d)) for d, k, n in zip(dates, strikes, notionals)]
    return storage, inventory, calls


def three_calls_valuation(num_sims, seed, forward_seed):
    val_date, rate, mean_reversion = Day(2019, 8, 29), 0.09, 16.5
    fwd, spot_vol = make_test_curves(val_date, Day(2020, 4, 1))
    storage, inventory, calls = three_calls(fwd)
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.BasisFunctions = ONE_FACTOR_BASIS
    b.ForwardCurve = fwd
    b.Inventory = inventory
    b.Storage = storage
    b.SettleDateRule = settle_rule
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(rate)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 100)
    b.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(mean_reversion, spot_vol), num_sims, seed, forward_seed)
    results = LsmcStorageValuation().Calculate(b.Build())
    expected, deltas = 0.0, {}
    for c in calls:
        iv = one_factor_implied_vol(val_date, c["expiry"], spot_vol, mean_reversion)
        value, delta = black76_call(val_date, fwd[c["expiry"]], iv, rate, c["strike"], c["expiry"], c["settle"])
        expected += value * c["notional"]
        deltas[c["expiry"]] = delta * c["notional"]
    return results, expected, deltas


def test_storage_like_three_call_options_npv_equals_black76():  # LsmcStorageValuationTest.cs:312-361
    results, expected, _ = three_calls_valuation(100_000, 11, 13)
    error = (results.Npv - expected) / expected
    two_standard_errors = 2.0 * results.ValuationSimStandardError / expected
    print(f"three calls: LSMC {results.Npv:.4f}  Black-76 {expected:.4f}  error {error:.5%}  2 s.e. {two_standard_errors:.5%}")
    assert -0.02 <= error <= two_standard_errors  # LSMC is a lower bound: not above the options but for sampling error


def test_storage_like_three_call_options_deltas_equal_black76_deltas_undiscounted():  # :363-423
    results, _, expected = three_calls_valuation(100_000, 11, 14)
    for day, delta in results.Deltas.items():
        if day in expected:
            assert abs(delta - expected[day]) / abs(expected[day]) <= 0.02, (str(day), delta, expected[day])
        else:
            assert delta == 0.0, str(day)


def reference_storage(kind):
    start, end = Day(2019, 8, 3), Day(2020, 4, 1)  # LsmcStorageValuationTest.cs:71-123
    b = CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
    if kind == "simple":
        b = b.WithConstantInjectWithdrawRange(-850.0, 625.0).WithZeroMinInventory().WithConstantMaxInventory(52_500.0)
    else:
        b = b.WithTimeAndInventoryVaryingInjectWithdrawRatesPiecewiseLinear([
            (start, [(0.0, (-702.7, 650.0)), (15_000.0, (-785.0, 552.5)), (30_000.0, (-790.6, 512.8)),
                     (40_000.0, (-825.6, 498.6)), (52_500.0, (-850.4, 480.0))]),
            (Day(2020, 2, 1), [(0.0, (-645.35, 650.0)), (13_000.0, (-656.0, 552.5)), (28_000.0, (-689.6, 512.8)),
                               (42_000.0, (-701.06, 498.6)), (52_500.0, (-718.04, 480.0))])])
    return (b.WithPerUnitInjectionCost(1.25, lambda d: d).WithNoCmdtyConsumedOnInject()
            .WithPerUnitWithdrawalCost(0.93, lambda d: d).WithNoCmdtyConsumedOnWithdraw()
            .WithNoCmdtyInventoryLoss().WithNoInventoryCost().MustBeEmptyAtEnd().Build())


def reference_builder(storage, basis, model, num_factors):
    val_date, end = Day(2019, 8, 29), Day(2020, 4, 1)  # :140-178
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = 5_685.0
    b.ForwardCurve = {d: 53.5 + math.sin(2.0 * math.pi / 365.0 * d.OffsetFrom(val_date)) * 24.6 for d in val_date.EnumerateTo(end)}
    b.Storage = storage
    b.SettleDateRule = lambda d: Month(d.Start.year, d.Start.month).Offset(1).First() + 19
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.055)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 100)
    b.BasisFunctions = basis
    b.SimulateWithMultiFactorModelAndMersenneTwister(model, 2_000, 11)
    params = b.Build()
    lsmc = LsmcStorageValuation().Calculate(params)
    intrinsic_npv, _ = oracle.intrinsic_calculate(FlatProblem(params, 2_000, num_factors))
    return lsmc, intrinsic_npv


@pytest.mark.parametrize("kind", ["simple", "ratchets"])
def test_one_factor_zero_mean_reversion_npv_approximately_equals_intrinsic(kind):  # :528-571
    lsmc, intrinsic = reference_builder(reference_storage(kind), ONE_FACTOR_BASIS, MultiFactorParameters.For1Factor(0.0, 0.95), 1)
    print(f"zero mean reversion, {kind}: LSMC {lsmc.Npv:.2f}  intrinsic {intrinsic:.2f}")
    assert abs(lsmc.Npv - intrinsic) / abs(intrinsic) <= 0.03


@pytest.mark.parametrize("kind", ["simple", "ratchets"])
def test_two_factor_very_low_vols_npv_approximately_equals_intrinsic(kind):  # :655-687
    model = MultiFactorParameters.For2Factors(0.61, (0.0, 0.001), (12.5, 0.001))
    basis = (BasisFunctionsBuilder.Ones + BasisFunctionsBuilder.AllMarkovFactorAllPositiveIntegerPowersUpTo(2, 2)
             + BasisFunctionsBuilder.SpotPricePower(1) + BasisFunctionsBuilder.SpotPricePower(2))  # :202-205
    lsmc, intrinsic = reference_builder(reference_storage(kind), basis, model, 2)
    assert abs(lsmc.Npv - intrinsic) / abs(intrinsic) <= 0.00001


def test_forced_inject_and_withdraw_npv_almost_equals_trivial_intrinsic_calc():  # :692-785, builder :788-851
    val_date, rate, mean_reversion = Day(2019, 8, 29), 0.09, 16.5
    start, end = Day(2019, 12, 1), Day(2020, 4, 1)
    inject_rate, inject_days, inject_start = 211.5, 20, Day(2019, 12, 20)
    withdraw_rate, withdraw_days, withdraw_start = 187.54, 15, Day(2020, 2, 5)
    inject_cost, inject_consumed, withdraw_cost, withdraw_consumed = 1.23, 0.01, 0.98, 0.015
    fwd, spot_vol = make_test_curves(val_date, end)

    def flat(rate_):
        return [(0.0, (rate_, rate_)), (10_000.0, (rate_, rate_))]
    sched = [(start, flat(0.0)), (inject_start, flat(inject_rate)), (inject_start + inject_days, flat(0.0)),
             (withdraw_start, flat(-withdraw_rate)), (withdraw_start + withdraw_days, flat(0.0))]
    storage = (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
               .WithTimeAndInventoryVaryingInjectWithdrawRatesPolynomial(sched)
               .WithPerUnitInjectionCost(inject_cost, lambda d: d + 10).WithFixedPercentCmdtyConsumedOnInject(inject_consumed)
               .WithPerUnitWithdrawalCost(withdraw_cost, lambda d: d + 4).WithFixedPercentCmdtyConsumedOnWithdraw(withdraw_consumed)
               .WithNoCmdtyInventoryLoss().WithNoInventoryCost().WithTerminalInventoryNpv(0.0).Build())
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = 0.0
    b.ForwardCurve = fwd
    b.Storage = storage
    b.SettleDateRule = settle_rule
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(rate)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 500)
    b.BasisFunctions = ONE_FACTOR_BASIS
    b.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(mean_reversion, spot_vol), 1_000, 11)
    results = LsmcStorageValuation().Calculate(b.Build())

    def df(day):
        return math.exp(-day.OffsetFrom(val_date) / 365.0 * rate)
    expected = 0.0
    for i in range(inject_days):
        d = inject_start + i
        expected += -fwd[d] * inject_rate * (1 + inject_consumed) * df(settle_rule(d)) - inject_cost * inject_rate * df(d + 10)
    for i in range(withdraw_days):
        d = withdraw_start + i
        expected += fwd[d] * withdraw_rate * (1 - withdraw_consumed) * df(settle_rule(d)) - withdraw_cost * withdraw_rate * df(d + 4)
    assert abs(results.Npv - expected) / abs(expected) <= 0.03
    assert np.isfinite(results.ValuationSimStandardError)


# ---------------------------------------------------------------------------------------------------------------------
# Trigger prices and returned simulation data on the reference's own fixture (LsmcStorageValuationTest.cs:180-192, 1131-1290)
# ---------------------------------------------------------------------------------------------------------------------
def reference_params(storage, num_sims, flags=None, inventory=5_685.0):
    val_date, end = Day(2019, 8, 29), Day(2020, 4, 1)  # :126-192
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = inventory
    b.ForwardCurve = {d: 53.5 + math.sin(2.0 * math.pi / 365.0 * d.OffsetFrom(val_date)) * 24.6 for d in val_date.EnumerateTo(end)}
    b.Storage = storage
    b.SettleDateRule = lambda d: Month(d.Start.year, d.Start.month).Offset(1).First() + 19
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.055)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 100)
    b.BasisFunctions = ONE_FACTOR_BASIS
    if flags is not None:
        b.SimulationDataReturned = flags
    b.SimulateWithMultiFactorModelAndMersenneTwister(MultiFactorParameters.For1Factor(12.5, 0.95), num_sims, 11)
    return b.Build()


def test_withdraw_trigger_price_higher_than_inject_trigger_price():  # :1212-1221 (500 simulations, :180)
    res = LsmcStorageValuation().Calculate(reference_params(reference_storage("simple"), 500))
    both = [t for t in res.TriggerPrices.values() if t.HasWithdrawPrice and t.HasInjectPrice]
    assert len(both) > 100
    for t in both:
        assert t.MaxWithdrawTriggerPrice > t.MaxInjectTriggerPrice


def test_trigger_volumes_consistent_with_max_inject_withdraw_rate():  # :1255-1290
    res = LsmcStorageValuation().Calculate(reference_params(reference_storage("simple"), 500))
    max_withdrawal_rate, max_injection_rate, max_inventory, min_inventory = 850.0, 625.0, 52_500.0, 0.0
    assert len(res.TriggerPrices) > 200
    for day, t in res.TriggerPrices.items():
        profile = res.TriggerPriceVolumeProfiles[day]
        expected_inventory = res.ExpectedStorageProfile[day].Inventory
        if t.HasInjectPrice:
            expected = min(max_injection_rate, max_inventory - expected_inventory)
            assert t.MaxInjectVolume == pytest.approx(expected, rel=1e-12, abs=1e-9)
            assert max(v for v, _ in profile.InjectTriggerPrices) == pytest.approx(expected, rel=1e-12, abs=1e-9)
        else:
            assert len(profile.InjectTriggerPrices) == 0
        if t.HasWithdrawPrice:
            expected = min(max_withdrawal_rate, expected_inventory - min_inventory)
            assert -t.MaxWithdrawVolume == pytest.approx(expected, rel=1e-12, abs=1e-9)
            assert -min(v for v, _ in profile.WithdrawTriggerPrices) == pytest.approx(expected, rel=1e-12, abs=1e-9)
        else:
            assert len(profile.WithdrawTriggerPrices) == 0


def test_no_withdraw_trigger_price_when_empty_and_no_inject_trigger_price_when_full():  # :1223-1253
    empty = LsmcStorageValuation().Calculate(reference_params(reference_storage("simple"), 500, inventory=0.0))
    first = empty.TriggerPrices[min(empty.TriggerPrices)]
    assert not first.HasWithdrawPrice and first.MaxWithdrawVolume is None and first.MaxWithdrawTriggerPrice is None
    assert len(empty.TriggerPriceVolumeProfiles[min(empty.TriggerPriceVolumeProfiles)].WithdrawTriggerPrices) == 0
    full = LsmcStorageValuation().Calculate(reference_params(reference_storage("simple"), 500, inventory=52_500.0))
    first = full.TriggerPrices[min(full.TriggerPrices)]
    assert not first.HasInjectPrice and first.MaxInjectVolume is None and first.MaxInjectTriggerPrice is None
    assert len(full.TriggerPriceVolumeProfiles[min(full.TriggerPriceVolumeProfiles)].InjectTriggerPrices) == 0


PANELS = ("RegressionSpotPriceSim", "ValuationSpotPriceSim", "InventoryBySim", "InjectWithdrawVolumeBySim", "CmdtyConsumedBySim",
          "InventoryLossBySim", "NetVolumeBySim", "PvByPeriodAndSim")


def test_simulation_data_returned_sizes():  # :1131-1192 (20 simulations)
    from storage_b200 import SimulationDataReturned as S
    num_sims = 20
    storage = reference_storage("simple")
    everything = LsmcStorageValuation().Calculate(reference_params(storage, num_sims, S.All))
    for name in PANELS:
        assert getattr(everything, name).shape[1] == num_sims, name
    assert len(everything.RegressionMarkovFactors) == 1 and len(everything.ValuationMarkovFactors) == 1
    for panel in everything.RegressionMarkovFactors + everything.ValuationMarkovFactors:
        assert panel.shape[1] == num_sims
    nothing = LsmcStorageValuation().Calculate(reference_params(storage, num_sims, S.Nothing))
    for name in PANELS:
        assert getattr(nothing, name).size == 0, name
    assert all(p.size == 0 for p in nothing.RegressionMarkovFactors + nothing.ValuationMarkovFactors)
    some = LsmcStorageValuation().Calculate(reference_params(storage, num_sims, S.Inventory | S.Pv | S.FactorsForValuation))
    for name in PANELS:
        expected_cols = num_sims if name in ("InventoryBySim", "PvByPeriodAndSim") else 0
        got = getattr(some, name)
        assert (got.shape[1] if got.size else 0) == expected_cols, name
    assert all(p.size == 0 for p in some.RegressionMarkovFactors)
    assert len(some.ValuationMarkovFactors) == 1 and some.ValuationMarkovFactors[0].shape[1] == num_sims
    assert some.Npv == everything.Npv == nothing.Npv
