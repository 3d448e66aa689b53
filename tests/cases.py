"""Synthetic valuation cases shaped like BASELINE.json's configs, with simulations generated on the host
by numpy so the SAME pre-simulated paths can be fed to the oracle and to the GPU through the caller-supplied-simulation
overload (LsmcValuationParameters.UseSpotSimResults, LsmcValuationParameters.cs:198-216)."""
import math
from types import SimpleNamespace

import numpy as np

from storage_b200 import (BasisFunctionsBuilder, CmdtyStorage, Day, FixedSpacingStateSpaceGridCalc, FlatProblem,
                          LsmcStorageValuation, LsmcValuationParameters, MultiFactorParameters, Sim,
                          SimulationDataReturned, SpotSimResultsFromPanels, StorageHelper,
                          settle_on_20th_of_following_month)


def forward_curve(val_date, end, base=53.5, amplitude=24.6):
    """F_t = 53.5 + 24.6 sin(2 pi t / 365) (tests/Cmdty.Storage.Test/Lsmc/LsmcStorageValuationTest.cs:156-162)."""
    return {d: base + math.sin(2.0 * math.pi / 365.0 * d.OffsetFrom(val_date)) * amplitude for d in val_date.EnumerateTo(end)}


def simulate_host(model, current, periods, fwd, num_sims, seed, antithetic=True):
    """numpy restatement of the multi-factor model (exact OU discretisation); returns spot [rows][N], factors F x [rows][N]."""
    rng = np.random.default_rng(seed)
    F = model.NumFactors
    mr = np.array([f[0] for f in model.factors])
    corr = np.asarray(model.corr, dtype=np.float64)
    rows = len(periods)
    spot = np.zeros((rows, num_sims))
    factors = np.zeros((F, rows, num_sims))
    x = np.zeros((F, num_sims))
    t_prev = 0.0

    def integral(a, t):
        return t if a == 0.0 else (1.0 - math.exp(-a * t)) / a
    for r, p in enumerate(periods):
        t = (p.First().ordinal - current.First().ordinal) / 365.0
        dt = t - t_prev
        cov = np.array([[corr[i, j] * integral(mr[i] + mr[j], dt) for j in range(F)] for i in range(F)])
        if dt > 0:
            chol = np.linalg.cholesky(cov)
            half = num_sims // 2 if antithetic else num_sims
            z = rng.standard_normal((F, half))
            if antithetic:
                z = np.concatenate([z, -z], axis=1)
                if z.shape[1] < num_sims:
                    z = np.concatenate([z, rng.standard_normal((F, num_sims - z.shape[1]))], axis=1)
            inc = chol @ z
        else:
            inc = np.zeros((F, num_sims))
        x = x * np.exp(-mr * dt)[:, None] + inc
        vol = np.array([model.vol(i, p) for i in range(F)])
        var = sum(vol[i] * vol[j] * corr[i, j] * integral(mr[i] + mr[j], t) for i in range(F) for j in range(F))
        spot[r] = fwd[p] * np.exp(vol @ x - 0.5 * var)
        factors[:, r, :] = x
        t_prev = t
    return spot, factors


def _finish(storage, val_date, inventory, model, basis, num_sims, seed, grid_points, flags=SimulationDataReturned.Nothing,
            extra_decisions=0, discount_deltas=False, tolerance=1e-10, rate=0.055):
    end = storage.EndPeriod
    fwd = forward_curve(val_date, end)
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = inventory
    b.ForwardCurve = fwd
    b.Storage = storage
    b.SettleDateRule = settle_on_20th_of_following_month
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(rate)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, grid_points)
    b.BasisFunctions = basis
    b.SimulationDataReturned = flags
    b.ExtraDecisions = extra_decisions
    b.DiscountDeltas = discount_deltas
    b.NumericalTolerance = tolerance
    sim_start = max(val_date.Offset(1), storage.StartPeriod)
    sim_periods = sim_start.EnumerateTo(end)
    reg_spot, reg_fac = simulate_host(model, val_date, sim_periods, fwd, num_sims, seed)
    val_spot, val_fac = simulate_host(model, val_date, sim_periods, fwd, num_sims, seed * 2 + 1)
    b.UseSpotSimResults(SpotSimResultsFromPanels(sim_periods, reg_spot, list(reg_fac)),
                        SpotSimResultsFromPanels(sim_periods, val_spot, list(val_fac)))
    params = b.Build()
    flat = FlatProblem(params, num_sims, model.NumFactors)
    return SimpleNamespace(params=params, builder=b, flat=flat, model=model, reg_spot=reg_spot, reg_factors=list(reg_fac),
                           val_spot=val_spot, val_factors=list(val_fac), num_sims=num_sims, sim_periods=sim_periods)


ONE_FACTOR_BASIS = (BasisFunctionsBuilder.Ones + BasisFunctionsBuilder.SpotPricePower(1) + BasisFunctionsBuilder.SpotPricePower(2)
                    + BasisFunctionsBuilder.SpotPricePower(3))
THREE_FACTOR_BASIS_TEXT = "1 + x0 + x0**2 + x1 + x2 + x0*x1 + s"  # BASELINE.json configs[1]


def simple_storage(start, end, terminal=None, loss=0.0):
    """The reference fixture's constant-rate storage (LsmcStorageValuationTest.cs:75-93)."""
    b = (CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
         .WithConstantInjectWithdrawRange(-850.0, 625.0)
         .WithZeroMinInventory()
         .WithConstantMaxInventory(52_500.0)
         .WithPerUnitInjectionCost(1.25, lambda d: d)
         .WithNoCmdtyConsumedOnInject()
         .WithPerUnitWithdrawalCost(0.93, lambda d: d)
         .WithNoCmdtyConsumedOnWithdraw())
    b = b.WithFixedPercentCmdtyInventoryLoss(loss) if loss else b.WithNoCmdtyInventoryLoss()
    b = b.WithNoInventoryCost()
    b = b.WithTerminalInventoryNpv(*terminal) if terminal else b.MustBeEmptyAtEnd()
    return b.Build()


def simple_one_factor(num_sims=2000, num_days=60, seed=11, val_offset=-5, inventory=5_685.0, grid_points=100, **kw):
    """BASELINE.json configs[0] shape: constant-parameter storage, 1-factor mean-reverting model (alpha 12.5, sigma 0.95),
    basis 1+s+s^2+s^3.  val_offset < 0: valuation date before the storage start; > 0: inside the active period."""
    start = Day(2019, 9, 1)
    end = start + num_days
    val_date = start + val_offset
    storage = simple_storage(start, end, terminal=kw.pop("terminal", None), loss=kw.pop("loss", 0.0))
    model = MultiFactorParameters.For1Factor(12.5, 0.95)
    basis = kw.pop("basis", None) or ONE_FACTOR_BASIS
    return _finish(storage, val_date, inventory, model, basis, num_sims, seed, grid_points, **kw)


RATCHETS_A = [(0.0, (-702.7, 650.0)), (15_000.0, (-785.0, 552.5)), (30_000.0, (-790.6, 512.8)), (40_000.0, (-825.6, 498.6)),
              (52_500.0, (-850.4, 480.0))]  # LsmcStorageValuationTest.cs:99-106
RATCHETS_B = [(0.0, (-645.35, 650.0)), (13_000.0, (-656.0, 552.5)), (28_000.0, (-689.6, 512.8)), (42_000.0, (-701.06, 498.6)),
              (52_500.0, (-718.04, 480.0))]  # LsmcStorageValuationTest.cs:107-114
OUTAGE = [(0.0, (0.0, 0.0)), (52_500.0, (0.0, 0.0))]


def ratchet_storage(start, end, outages=True, kind="linear"):
    """BASELINE.json configs[2]: time- and inventory-varying ratchets with outages, inventory loss, commodity consumed,
    inventory cost, empty at end."""
    n = end.OffsetFrom(start)
    change = start + max(2, (2 * n) // 3)
    sched = [(start, RATCHETS_A)]
    if outages and n >= 30:
        o1 = start + n // 4
        sched += [(o1, OUTAGE), (o1 + min(10, n // 8), RATCHETS_A)]
    sched.append((change, RATCHETS_B))
    if outages and n >= 30:
        o2 = change + max(2, n // 10)
        if o2 + 6 < end:
            sched += [(o2, OUTAGE), (o2 + min(10, n // 8), RATCHETS_B)]
    b = CmdtyStorage.Builder.WithActiveTimePeriod(start, end)
    if kind == "linear":
        b = b.WithTimeAndInventoryVaryingInjectWithdrawRatesPiecewiseLinear(sched)
    elif kind == "step":
        def stepify(rows):
            return rows[:-1] + [(rows[-1][0], rows[-2][1])] if rows is not OUTAGE else rows
        b = b.WithStepRatchets([(d, stepify(r)) for d, r in sched])
    else:
        b = b.WithTimeAndInventoryVaryingInjectWithdrawRatesPolynomial(sched)
    return (b.WithPerUnitInjectionCost(1.25, lambda d: d)
            .WithFixedPercentCmdtyConsumedOnInject(0.01)
            .WithPerUnitWithdrawalCost(0.93, lambda d: d)
            .WithFixedPercentCmdtyConsumedOnWithdraw(0.015)
            .WithFixedPercentCmdtyInventoryLoss(0.001)
            .WithFixedPerUnitInventoryCost(0.002)
            .MustBeEmptyAtEnd()
            .Build())


def three_factor_model(start, end):
    """README three-factor seasonal parameters (samples/csharp/Cmdty.Storage.Samples.LsmcThreeFactor/Program.cs:88-95)."""
    return MultiFactorParameters.For3FactorSeasonal(90.6, 0.7, 0.17, 0.32, start, end)


def three_factor_ratchets(num_sims=2000, num_days=60, seed=11, val_offset=-1, inventory=5_250.0, grid_points=100, kind="linear",
                          outages=True, **kw):
    start = Day(2019, 9, 1)
    end = start + num_days
    val_date = start + val_offset
    storage = ratchet_storage(start, end, outages=outages, kind=kind)
    model = three_factor_model(start, end)
    basis = kw.pop("basis", None) or (BasisFunctionsBuilder.Ones + Sim.X0 + Sim.X0.Pow(2) + Sim.X1 + Sim.X2 + Sim.X0 * Sim.X1 + Sim.S)
    return _finish(storage, val_date, inventory, model, basis, num_sims, seed, grid_points, **kw)


def run_gpu(case, keep_decisions=True, device=0, reference_operation_order=False):
    return LsmcStorageValuation(device=device).Calculate(case.params, keep_decisions=keep_decisions,
                                                        reference_operation_order=reference_operation_order)
