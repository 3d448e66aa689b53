"""The reference's own LSMC regression fixtures (src/Cmdty.Storage.Python/tests/test_multi_factor.py), rebuilt through this
repo's host API: same storage, curves, model, basis, seeds.  The simulated paths come from oracle/reference_sim.py, the
restatement of the reference's un-vendored simulation chain that reproduces its golden simulator values digit for digit."""
import csv
import os

import numpy as np

from oracle import reference_sim
from storage_b200 import (BasisFunctionsBuilder, CmdtyStorage, Day, FixedSpacingStateSpaceGridCalc, FlatProblem,
                          LsmcValuationParameters, MultiFactorParameters, SimulationDataReturned, SpotSimResultsFromPanels,
                          StorageHelper, settle_on_20th_of_following_month)

GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")

# literals asserted by the reference (assertAlmostEqual places=6 for npv, assertEqual for the standard error)
GOLDENS = {
    "multi_factor_test-1": dict(npv=1780380.7581833513, std_error=21405.33891571013, lines="test_multi_factor.py:99,103"),
    "three_factor_test-1": dict(npv=1766460.137569665, std_error=18459.69670786117, lines="test_multi_factor.py:259,263"),
}


class ReferenceCase:
    pass


def build(name, num_sims=500, seed=11, flags=SimulationDataReturned.Nothing):
    """test_multi_factor.py:34-97 (two factors) / :199-258 (three-factor seasonal): storage, curve, model, simulation."""
    start, end = Day(2019, 12, 1), Day(2020, 4, 1)
    storage = (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
               .WithConstantInjectWithdrawRange(-700.0, 700.0).WithConstantMinInventory(0.0).WithConstantMaxInventory(100000.0)
               .WithPerUnitInjectionCost(1.23).WithNoCmdtyConsumedOnInject().WithPerUnitWithdrawalCost(0.98).WithNoCmdtyConsumedOnWithdraw()
               .WithNoCmdtyInventoryLoss().WithNoInventoryCost().MustBeEmptyAtEnd().Build())
    val_date, switch = Day(2019, 8, 29), Day(2020, 3, 12)
    fwd = {d: (23.87 if d < switch else 150.32) for d in val_date.EnumerateTo(end)}
    if name == "multi_factor_test-1":
        model = MultiFactorParameters.For2Factors(0.64, (0.0, 0.14), (16.2, 1.15))  # factors (long-term, spot), correlation 0.64
        basis = BasisFunctionsBuilder.Parse("1 + x0 + x0**2 + x1 + x1*x1", 2)
    else:
        model = MultiFactorParameters.For3FactorSeasonal(16.2, 1.15, 0.14, 0.18, val_date, end)
        # '1 + x_st + x_sw + x_lt + x_st**2 + x_sw**2 + x_lt**2' with x_st -> x0, x_lt -> x1, x_sw -> x2 (multi_factor.py:125-126)
        basis = BasisFunctionsBuilder.Parse("1 + x0 + x2 + x1 + x0**2 + x2**2 + x1**2", 3)
    F = model.NumFactors
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = 0.0
    b.ForwardCurve = fwd
    b.Storage = storage
    b.SettleDateRule = settle_on_20th_of_following_month
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.03)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, 100)
    b.NumericalTolerance = 1e-12  # multi_factor.py:116
    b.BasisFunctions = basis
    b.SimulationDataReturned = flags
    b.SimulateWithMultiFactorModelAndMersenneTwister(model, num_sims, seed, seed)  # placeholder, replaced below
    probe = FlatProblem(b.Build(), num_sims, F)
    periods = probe.sim_periods
    times = [(p.First().ordinal - val_date.First().ordinal) / 365.0 for p in periods]  # TimeFunctions.Act365 from currentPeriod.Start
    vol = np.array([[model.vol(i, p) for p in periods] for i in range(F)])
    # SimulateWithMultiFactorModelAndMersenneTwister(model, numSims, seed, fwd_sim_seed = seed): antithetic, and the valuation
    # generator is a second MersenneTwisterGenerator with the same seed, i.e. the same paths (LVP.cs:185-196)
    spot, factors = reference_sim.simulate([f[0] for f in model.factors], vol, model.corr, times, [fwd[p] for p in periods], num_sims, seed, True)
    b.UseSpotSimResults(SpotSimResultsFromPanels(periods, spot, list(factors)), SpotSimResultsFromPanels(periods, spot, list(factors)))
    case = ReferenceCase()
    case.name, case.builder, case.params = name, b, b.Build()
    case.flat = FlatProblem(case.params, num_sims, F)
    case.spot, case.factors, case.num_sims = spot, list(factors), num_sims
    case.golden = GOLDENS[name]
    return case


def load_csv(name, which):
    with open(os.path.join(GOLDEN_DIR, f"reference_{name}_{which}.csv"), newline="") as f:
        rows = list(csv.reader(f))
    if which != "deltas":
        rows = rows[1:]
    return np.array([[float(x) if x != "" else np.nan for x in r[1:]] for r in rows])


def compare_with_csvs(name, deltas, profile, trigger, rel):
    """deltas [P+1]; profile [P+1][5] (inventory, volume, consumed, loss, period pv); trigger: per period (has_inject, max inject
    volume, price, has_withdraw, max withdraw volume, price).  Returns the largest relative differences found."""
    g_deltas = load_csv(name, "deltas")[:, 0]
    g_profile = load_csv(name, "expected_profile")  # inventory, inject_withdraw_volume, cmdty_consumed, inventory_loss, net_volume, period_pv
    g_trigger = load_csv(name, "trigger_prices")
    worst = {}
    assert len(g_deltas) == len(deltas)
    worst["deltas"] = float(np.abs(deltas - g_deltas).max() / np.abs(g_deltas).max())
    ours = np.column_stack([profile[:, 0], profile[:, 1], profile[:, 2], profile[:, 3], -profile[:, 1] - profile[:, 2], profile[:, 4]])
    assert ours.shape == g_profile.shape
    worst["expected_profile"] = max(float(np.abs(ours[:, c] - g_profile[:, c]).max() / max(np.abs(g_profile[:, c]).max(), 1e-300))
                                    for c in range(6))
    assert len(trigger) == len(g_trigger) - 0 or len(trigger) == len(g_trigger)
    diffs = [0.0]
    for row, t in zip(g_trigger, trigger):
        has_inj, inj_vol, inj_price, has_wd, wd_vol, wd_price = t
        assert bool(has_inj) == (not np.isnan(row[0])) and bool(has_wd) == (not np.isnan(row[2]))
        if has_inj:
            diffs += [abs(inj_vol - row[0]) / abs(row[0]), abs(inj_price - row[1]) / abs(row[1])]
        if has_wd:
            diffs += [abs(wd_vol - row[2]) / abs(row[2]), abs(wd_price - row[3]) / abs(row[3])]
    worst["trigger_prices"] = float(max(diffs))
    for k, v in worst.items():
        assert v <= rel, f"{name} {k}: {v:.3e} > {rel:g}"
    return worst
