"""Closed-form and invariance tests of Calculate ported from the reference's LsmcStorageValuationTest.cs
(tests/Cmdty.Storage.Test/Lsmc/): they need no particular random stream, so they run against the
CPU oracle here (no GPU) and against the CUDA library on the GPU box (-m gpu)."""
import threading

import numpy as np
import pytest

from oracle import oracle
from storage_b200 import (ArgumentException, BasisFunctionsBuilder, CmdtyStorage, Day, InvalidOperationException,
                          InventoryConstraintsCannotBeFulfilledException, LsmcStorageValuation, LsmcValuationParameters,
                          OperationCanceledException, SimulationDataReturned, UnsupportedStorageFeature)
from tests import cases


def run_oracle(case, mode=1):
    return oracle.lsmc_calculate(case.flat, case.reg_spot, case.reg_factors, case.val_spot, case.val_factors, mode=mode)


class OracleRunner:
    name = "oracle"

    def run(self, case):
        r = run_oracle(case)
        return dict(npv=r.npv, deltas=r.deltas, profile=r.profile, volumes=r.inject_withdraw_volume, trigger=r.trigger,
                    pv_by_sim=r.pv_by_sim)


class GpuRunner:
    name = "gpu"

    def run(self, case):
        case.params.SimulationDataReturned = SimulationDataReturned.All
        g = cases.run_gpu(case)
        periods = case.flat.periods
        prof = np.array([[s.Inventory, s.InjectWithdrawVolume, s.CmdtyConsumed, s.InventoryLoss, s.PeriodPv]
                         for s in (g.ExpectedStorageProfile[p] for p in periods)])
        return dict(npv=g.Npv, deltas=np.array([g.Deltas[p] for p in periods]), profile=prof, volumes=g.InjectWithdrawVolumeBySim,
                    trigger=[g.TriggerPrices[p] for p in periods[:-1]], pv_by_sim=g.PvBySim)


@pytest.fixture(params=["oracle", pytest.param("gpu", marks=pytest.mark.gpu)])
def runner(request):
    return OracleRunner() if request.param == "oracle" else GpuRunner()


# ---------------------------------------------------------------- at / after the end of the storage (host-side guards)
def _builder_at(current, storage, inventory=0.0):
    case = cases.simple_one_factor(num_sims=16, num_days=10, inventory=100.0)
    b = case.builder.Clone()
    b._sim_spec = case.builder._sim_spec
    b.Storage = storage
    b.CurrentPeriod = current
    b.Inventory = inventory
    b.ForwardCurve = cases.forward_curve(storage.StartPeriod - 10, storage.EndPeriod + 5)
    return b


def test_current_period_after_storage_end_gives_zero_npv_and_empty_deltas():  # LsmcStorageValuationTest.cs:209-229
    st = cases.simple_storage(Day(2019, 9, 1), Day(2019, 9, 11))
    res = LsmcStorageValuation.WithNoLogger.Calculate(_builder_at(st.EndPeriod + 1, st).Build())
    assert res.Npv == 0.0 and res.Deltas == {}


def test_current_period_equal_to_end_must_be_empty():  # :232-254, LSV.cs:73-81
    st = cases.simple_storage(Day(2019, 9, 1), Day(2019, 9, 11))
    res = LsmcStorageValuation.WithNoLogger.Calculate(_builder_at(st.EndPeriod, st, 0.0).Build())
    assert res.Npv == 0.0 and res.Deltas == {}
    with pytest.raises(InventoryConstraintsCannotBeFulfilledException):
        LsmcStorageValuation.WithNoLogger.Calculate(_builder_at(st.EndPeriod, st, 1.0).Build())


def test_current_period_equal_to_end_with_terminal_value():  # :258-272
    st = cases.simple_storage(Day(2019, 9, 1), Day(2019, 9, 11), terminal=(-999.0, 0.0, 0.0, 1.0))
    params = _builder_at(st.EndPeriod, st, 120.5).Build()
    res = LsmcStorageValuation.WithNoLogger.Calculate(params)
    assert res.Npv == params.ForwardCurve[st.EndPeriod] * 120.5 - 999.0


def test_negative_inventory_and_forward_curve_checks():  # LSV.cs:64-65, 94-98
    st = cases.simple_storage(Day(2019, 9, 1), Day(2019, 9, 11))
    with pytest.raises(ArgumentException):
        LsmcStorageValuation.WithNoLogger.Calculate(_builder_at(Day(2019, 8, 30), st, -1.0).Build())
    b = _builder_at(Day(2019, 8, 30), st, 10.0)
    b.ForwardCurve = cases.forward_curve(Day(2019, 8, 30), Day(2019, 9, 5))
    with pytest.raises(ArgumentException, match="does not extend"):
        LsmcStorageValuation.WithNoLogger.Calculate(b.Build())
    b.ForwardCurve = cases.forward_curve(Day(2019, 9, 3), Day(2019, 9, 15))
    with pytest.raises(ArgumentException, match="starts too late"):
        LsmcStorageValuation.WithNoLogger.Calculate(b.Build())


def test_builder_reports_unset_members_and_negative_extra_decisions():  # LVP.cs:124-144
    b = LsmcValuationParameters.Builder()
    with pytest.raises(InvalidOperationException, match="CurrentPeriod"):
        b.Build()
    case = cases.simple_one_factor(num_sims=16, num_days=10, inventory=100.0)
    c = case.builder.Clone()
    c._sim_spec = case.builder._sim_spec
    c.ExtraDecisions = -1
    with pytest.raises(InvalidOperationException, match="ExtraDecisions"):
        c.Build()
    c.ExtraDecisions = 0
    c.Inventory = None
    with pytest.raises(InvalidOperationException, match="Inventory"):
        c.Build()


def test_simulations_must_cover_the_storage_periods():  # CheckSpotSim, LVP.cs:219-231
    from storage_b200 import SpotSimResultsFromPanels
    case = cases.simple_one_factor(num_sims=16, num_days=10, inventory=100.0)
    short = SpotSimResultsFromPanels(case.sim_periods[:-2], case.reg_spot[:-2], [f[:-2] for f in case.reg_factors])
    b = case.builder.Clone()
    b.UseSpotSimResults(short, short)
    with pytest.raises(ArgumentException, match="does not contain simulations for period"):
        LsmcStorageValuation.WithNoLogger.Calculate(b.Build())


def test_opaque_func_overloads_are_rejected_not_approximated():  # SURVEY hard part 5
    b = CmdtyStorage.Builder.WithActiveTimePeriod(Day(2020, 1, 1), Day(2020, 1, 5)).WithConstantInjectWithdrawRange(-1, 1)
    with pytest.raises(UnsupportedStorageFeature):
        b.WithInjectionCost(lambda p, inv, vol: [])
    with pytest.raises(UnsupportedStorageFeature):
        b.WithTerminalInventoryNpv(lambda price, inv: 0.0)


# ---------------------------------------------------------------- closed forms that need a run
def test_day_before_end_npv_and_delta(runner):  # LsmcStorageValuationTest.cs:276-306
    case = cases.simple_one_factor(num_sims=200, num_days=10, val_offset=9, inventory=352.14)
    out = runner.run(case)
    p = case.params
    spot = p.ForwardCurve[p.CurrentPeriod]
    df = p.DiscountFactors(p.CurrentPeriod, p.SettleDateRule(p.CurrentPeriod))
    assert out["npv"] == pytest.approx(352.14 * spot * df - 0.93 * 352.14, abs=1e-8)
    assert out["deltas"][0] == pytest.approx(352.14, abs=1e-10)


def _forced_storage(start, end, forced_inject, forced_withdraw_from):
    """Forced injection at the start, forced withdrawal at the end (LsmcStorageValuationTest.cs:793-820 shape)."""
    ranges = {}
    for d in start.EnumerateTo(end):
        if d < forced_withdraw_from:
            ranges[d] = (forced_inject, forced_inject)
        else:
            ranges[d] = (-forced_inject * 1.0, -forced_inject * 1.0)
    from storage_b200 import InjectWithdrawRange
    return (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
            .WithInjectWithdrawRangeSeries({d: InjectWithdrawRange(*r) for d, r in ranges.items()})
            .WithZeroMinInventory().WithConstantMaxInventory(10_000.0)
            .WithPerUnitInjectionCost(1.25).WithNoCmdtyConsumedOnInject()
            .WithPerUnitWithdrawalCost(0.93).WithNoCmdtyConsumedOnWithdraw()
            .WithNoCmdtyInventoryLoss().WithNoInventoryCost().MustBeEmptyAtEnd().Build())


def test_forced_inject_then_withdraw_volumes_equal_forced_rates(runner):  # :863-919
    start = Day(2019, 9, 1)
    end = start + 20
    st = _forced_storage(start, end, 210.5, start + 10)
    model = cases.MultiFactorParameters.For1Factor(12.5, 0.95)
    case = cases._finish(st, start - 3, 0.0, model, cases.ONE_FACTOR_BASIS, 64, 5, 50)
    out = runner.run(case)
    vol = out["volumes"]
    np.testing.assert_allclose(vol[:10], 210.5, rtol=0, atol=1e-12)
    np.testing.assert_allclose(vol[10:20], -210.5, rtol=0, atol=1e-12)
    fwd, p = case.params.ForwardCurve, case.params
    expected = 0.0
    for i, d in enumerate(case.flat.periods[:-1]):
        v = 210.5 if i < 10 else -210.5
        df = p.DiscountFactors(p.CurrentPeriod, p.SettleDateRule(d))
        cost_df = p.DiscountFactors(p.CurrentPeriod, d)
        mean_spot = case.val_spot[i - case.flat.first_sim].mean()  # the decisions are forced, so the NPV is linear in the sample means
        expected += -v * mean_spot * df - (1.25 if v > 0 else 0.93) * abs(v) * cost_df
    assert out["npv"] == pytest.approx(expected, rel=1e-9)
    assert fwd[case.flat.periods[0]] > 0


def test_discounted_deltas_ratio(runner):  # :921-946
    plain = cases.simple_one_factor(num_sims=300, num_days=25, seed=3)
    disc = cases.simple_one_factor(num_sims=300, num_days=25, seed=3, discount_deltas=True)
    a, b = runner.run(plain), runner.run(disc)
    p = plain.params
    for i, d in enumerate(plain.flat.periods[:-1]):
        df = p.DiscountFactors(p.CurrentPeriod, p.SettleDateRule(d))
        assert b["deltas"][i] == pytest.approx(a["deltas"][i] * df, rel=1e-14, abs=1e-12)
    assert a["npv"] == b["npv"]


def test_no_trigger_prices_at_empty_or_full(runner):  # :1223-1253
    case = cases.simple_one_factor(num_sims=200, num_days=15, inventory=0.0, val_offset=-2)
    t0 = runner.run(case)["trigger"][0]
    has_withdraw = t0.HasWithdrawPrice if hasattr(t0, "HasWithdrawPrice") else bool(t0.has_withdraw_price)
    has_inject = t0.HasInjectPrice if hasattr(t0, "HasInjectPrice") else bool(t0.has_inject_price)
    assert not has_withdraw and has_inject  # empty storage: nothing to withdraw


def test_terminal_value_quirk_switch(runner):
    """with !MustBeEmptyAtEnd the reference adds the RUNNING SUM of terminal NPVs to each path (LSV.cs:606-609);
    reference_compat_terminal_pv=0 gives every path its own terminal NPV instead."""
    compat = cases.simple_one_factor(num_sims=50, num_days=12, seed=2, terminal=(-999.0, 0.0, 0.0, 1.0))
    own = cases.simple_one_factor(num_sims=50, num_days=12, seed=2, terminal=(-999.0, 0.0, 0.0, 1.0))
    own.params.ReferenceCompatTerminalPv = False
    own.flat.struct.reference_compat_terminal_pv = 0
    a, b = runner.run(compat), runner.run(own)
    assert a["profile"][-1, 4] == pytest.approx(b["profile"][-1, 4], rel=1e-12)  # mean terminal PV is the same number
    assert abs(a["npv"]) > 5 * abs(b["npv"])  # ...but the compat NPV carries the cumulative sum


# ---------------------------------------------------------------- GPU-only behaviour of the call
@pytest.mark.gpu
def test_progress_is_ascending_and_ends_at_one():  # :948-972, SURVEY Q10: 2P+1 notifications
    case = cases.simple_one_factor(num_sims=200, num_days=12)
    seen = []
    case.params.OnProgressUpdate = seen.append
    cases.run_gpu(case)
    P = case.flat.num_periods
    assert len(seen) == 2 * P + 1
    assert all(b >= a for a, b in zip(seen, seen[1:])) and seen[-1] == 1.0
    assert seen[P - 1] == pytest.approx(0.66, abs=1e-9)


@pytest.mark.gpu
def test_cancellation_raises():  # :974-993
    case = cases.simple_one_factor(num_sims=200, num_days=12)
    token = threading.Event()
    token.set()
    case.params.CancellationToken = token
    with pytest.raises(OperationCanceledException):
        cases.run_gpu(case)


@pytest.mark.gpu
def test_simulation_data_returned_flags_do_not_change_results():  # :995-1129
    base = cases.run_gpu(cases.simple_one_factor(num_sims=300, num_days=15, seed=9))
    for flags in (SimulationDataReturned.Inventory, SimulationDataReturned.Pv | SimulationDataReturned.NetVolume,
                  SimulationDataReturned.All):
        case = cases.simple_one_factor(num_sims=300, num_days=15, seed=9)
        case.params.SimulationDataReturned = flags
        res = cases.run_gpu(case)
        assert res.Npv == base.Npv and res.ValuationSimStandardError == base.ValuationSimStandardError
        assert list(res.Deltas.values()) == list(base.Deltas.values())
        P = case.flat.num_periods
        assert res.InventoryBySim.shape == ((P + 1, 300) if flags & SimulationDataReturned.Inventory else (0, 0))
        assert res.PvByPeriodAndSim.shape == ((P + 1, 300) if flags & SimulationDataReturned.Pv else (0, 0))
        if flags & SimulationDataReturned.SpotPricesForValuation:
            np.testing.assert_array_equal(res.ValuationSpotPriceSim, case.val_spot)
            np.testing.assert_array_equal(res.RegressionMarkovFactors[0], case.reg_factors[0])


@pytest.mark.gpu
def test_runs_are_bit_reproducible():
    a = cases.run_gpu(cases.three_factor_ratchets(num_sims=2000, num_days=30, seed=5))
    b = cases.run_gpu(cases.three_factor_ratchets(num_sims=2000, num_days=30, seed=5))
    assert a.Npv == b.Npv and a.ValuationSimStandardError == b.ValuationSimStandardError
    np.testing.assert_array_equal(a.DecisionIndexBySim, b.DecisionIndexBySim)


@pytest.mark.gpu
def test_basis_expression_parser_matches_operator_form():  # BasisFunctionsBuilderTest.cs:39-114
    from storage_b200 import Sim
    parsed = BasisFunctionsBuilder.Parse("1 + s + s*s + s*s**3 + x0 + x1**3 + s*x2**4", 3)
    built = (BasisFunctionsBuilder.Ones + Sim.Spot + Sim.Spot * Sim.Spot + Sim.Spot * Sim.Spot.Pow(3) + Sim.X0 + Sim.X1.Pow(3)
             + Sim.S * Sim.X2.Pow(4))
    assert [f.key() for f in parsed] == [f.key() for f in built]
