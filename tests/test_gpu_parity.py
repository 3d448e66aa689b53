"""GPU parity: the CUDA path (through the C ABI / host API mirror) against the CPU oracle on identical pre-simulated paths
fed through the caller-supplied-simulation overload (north star).

Tolerances (BASELINE.json north_star): NPV, deltas, profile within 1e-9 relative in FP64; argmax decisions identical.
"Identical decisions" is checked on decision *volumes*: when the inventory-space bound makes two entries of a decision set
coincide to ~1e-12 (e.g. max-rate withdrawal vs withdrawal constrained by the next step's maximum inventory) the index of
the maximum is decided by the last ulp of the continuation value, so index flips between numerically identical decisions
are reported but not counted as mismatches.
"""
import numpy as np
import pytest

from oracle import oracle
from storage_b200 import SimulationDataReturned
from tests import cases

pytestmark = pytest.mark.gpu

REL = 1e-9


def material_mismatches(gpu, ref):
    idx_diff = gpu.DecisionIndexBySim != ref.decision_index
    vol_diff = np.abs(gpu.InjectWithdrawVolumeBySim[:-1] - ref.inject_withdraw_volume[:-1]) > 1e-6
    return int(np.count_nonzero(idx_diff)), int(np.count_nonzero(vol_diff))


def check_case(case, rel=REL):
    """Both arithmetic forms of the backward kernel (default fused payoff, and the reference's operation order) against the oracle."""
    case.params.SimulationDataReturned = SimulationDataReturned.All
    ref = oracle.lsmc_calculate(case.flat, case.reg_spot, case.reg_factors, case.val_spot, case.val_factors, mode=1)
    for reference_order in (False, True):
        gpu = cases.run_gpu(case, reference_operation_order=reference_order)
        compare_with_oracle(case, gpu, ref, rel)
    return gpu, ref


def compare_with_oracle(case, gpu, ref, rel):
    raw, material = material_mismatches(gpu, ref)
    assert material == 0, f"{material} material decision mismatches ({raw} raw index flips)"
    assert gpu.Npv == pytest.approx(ref.npv, rel=rel)
    assert gpu.ValuationSimStandardError == pytest.approx(ref.std_error, rel=1e-8)
    assert gpu.BackwardNpv == pytest.approx(ref.backward_npv, rel=rel)
    P = case.flat.num_periods
    periods = case.flat.periods
    deltas = np.array([gpu.Deltas[p] for p in periods])
    scale = np.abs(ref.deltas).max() + 1e-300
    assert np.abs(deltas - ref.deltas).max() <= rel * scale
    prof = np.array([[sp.Inventory, sp.InjectWithdrawVolume, sp.CmdtyConsumed, sp.InventoryLoss, sp.PeriodPv]
                     for sp in (gpu.ExpectedStorageProfile[p] for p in periods)])
    for col in range(5):
        s = np.abs(ref.profile[:, col]).max() + 1e-300
        assert np.abs(prof[:, col] - ref.profile[:, col]).max() <= rel * s, f"profile column {col}"
    # per-sim panels
    np.testing.assert_allclose(gpu.PvBySim, ref.pv_by_sim, rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(gpu.InventoryBySim, ref.inventory, rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(gpu.InjectWithdrawVolumeBySim, ref.inject_withdraw_volume, rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(gpu.CmdtyConsumedBySim, ref.cmdty_consumed, rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(gpu.InventoryLossBySim, ref.inventory_loss, rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(gpu.NetVolumeBySim, ref.net_volume, rtol=1e-9, atol=1e-6)
    np.testing.assert_allclose(gpu.PvByPeriodAndSim, ref.pv, rtol=1e-9, atol=1e-5)
    # trigger prices
    for i in range(P):
        t_gpu, t_ref = gpu.TriggerPriceVolumeProfiles[periods[i]], ref.trigger[i]
        assert len(t_gpu.InjectTriggerPrices) == t_ref.num_inject and len(t_gpu.WithdrawTriggerPrices) == t_ref.num_withdraw
        for j, (vol, price) in enumerate(t_gpu.InjectTriggerPrices):
            assert vol == pytest.approx(t_ref.inject_volume[j], rel=1e-9, abs=1e-9)
            assert price == pytest.approx(t_ref.inject_price[j], rel=1e-7, abs=1e-7)
        for j, (vol, price) in enumerate(t_gpu.WithdrawTriggerPrices):
            assert vol == pytest.approx(t_ref.withdraw_volume[j], rel=1e-9, abs=1e-9)
            assert price == pytest.approx(t_ref.withdraw_price[j], rel=1e-7, abs=1e-7)
    return gpu, ref


def test_config0_one_factor_cubic_basis():
    """BASELINE.json configs[0] shape (2,000 paths, basis 1+s+s^2+s^3), 60 daily steps."""
    check_case(cases.simple_one_factor(num_sims=2000, num_days=60, seed=11))


def test_config0_full_year():
    """configs[0] at its full size: 365 daily steps, 2,000 paths, 100-point grid."""
    check_case(cases.simple_one_factor(num_sims=2000, num_days=365, seed=7, val_offset=-1))


def test_valuation_date_inside_active_period():
    """CurrentPeriod >= StartPeriod: period 0 uses the forward price and the sample-mean continuation (LSV.cs:174-185)."""
    check_case(cases.simple_one_factor(num_sims=1500, num_days=45, seed=5, val_offset=12))


def test_config1_three_factor_basis():
    """configs[1] shape: 3-factor seasonal model, basis 1+x0+x0^2+x1+x2+x0*x1+s, piecewise-linear ratchets."""
    check_case(cases.three_factor_ratchets(num_sims=3000, num_days=90, seed=11, outages=False))


def test_config2_ratchets_outages_loss_costs():
    """configs[2] shape: time- and inventory-varying ratchets, outages, inventory loss, consumed commodity, inventory cost."""
    check_case(cases.three_factor_ratchets(num_sims=2000, num_days=120, seed=3))


@pytest.mark.parametrize("kind", ["step", "poly"])
def test_other_constraint_kinds(kind):
    check_case(cases.three_factor_ratchets(num_sims=1000, num_days=50, seed=9, kind=kind))


def test_terminal_inventory_value_and_cumulative_quirk():
    """!MustBeEmptyAtEnd with TerminalStorageNpv = price*inventory - 999 (LsmcStorageValuationTest.cs:125), including the
    reference's running-sum terminal PV (LSV.cs:606-609)."""
    check_case(cases.simple_one_factor(num_sims=1000, num_days=40, seed=2, terminal=(-999.0, 0.0, 0.0, 1.0)))


def test_extra_decisions_and_discounted_deltas():
    check_case(cases.simple_one_factor(num_sims=1000, num_days=40, seed=4, extra_decisions=2, discount_deltas=True))


def test_ragged_path_count():
    """N not a multiple of the tile (odd, prime): tail handling of the double2 tiles."""
    check_case(cases.simple_one_factor(num_sims=1009, num_days=30, seed=6))
    check_case(cases.simple_one_factor(num_sims=7, num_days=10, seed=6, grid_points=20))


def test_rank_deficient_design_matrix_is_reported():
    """Fewer paths than basis functions: the reference's QR would produce non-finite coefficients; the library reports it."""
    from storage_b200 import NumericalError
    with pytest.raises(NumericalError):
        cases.run_gpu(cases.simple_one_factor(num_sims=3, num_days=10, seed=6, grid_points=20))


def test_single_period_storage():
    check_case(cases.simple_one_factor(num_sims=500, num_days=1, seed=8, inventory=300.0))


def test_larger_path_count_config1():
    """configs[1] nearer its size: 20,000 paths (the oracle's exact mode needs ~10 s here)."""
    check_case(cases.three_factor_ratchets(num_sims=20_000, num_days=60, seed=21, outages=False))


@pytest.mark.parametrize("num_basis", [1, 2, 3])
def test_narrow_bases(num_basis):
    """Bases of one to three monomials (the constant alone regresses on the sample mean of every period)."""
    from storage_b200 import BasisFunctionsBuilder
    basis = BasisFunctionsBuilder.Ones
    for power in range(1, num_basis):
        basis = basis + BasisFunctionsBuilder.SpotPricePower(power)
    check_case(cases.simple_one_factor(num_sims=1500, num_days=25, seed=23, basis=basis))


@pytest.mark.parametrize("variant", ["row_ring", "direct_sweep", "general"])
@pytest.mark.parametrize("num_basis", [10, 13])
def test_wide_bases(monkeypatch, variant, num_basis):
    """Bases of 10 and 13 monomials: the kernels' 12- and 16-wide instantiations (two 8-row blocks of regression sums per grid
    block, 16-wide partial-sum rows), in all three backward kernels."""
    from storage_b200 import BasisFunctionsBuilder, Sim
    if variant == "direct_sweep":
        monkeypatch.setenv("LSMC_SWEEP_RING", "0")
    elif variant == "general":
        monkeypatch.setenv("LSMC_SWEEP", "0")
    basis = (BasisFunctionsBuilder.Ones + Sim.X0 + Sim.X0.Pow(2) + Sim.X1 + Sim.X2 + Sim.X0 * Sim.X1 + Sim.S + Sim.S.Pow(2)
             + Sim.X1.Pow(2) + Sim.X2.Pow(2))
    if num_basis == 13:
        basis = basis + Sim.X0 * Sim.X2 + Sim.X1 * Sim.X2 + Sim.S * Sim.X0
    check_case(cases.three_factor_ratchets(num_sims=3000, num_days=30, seed=17, basis=basis))


@pytest.mark.parametrize("gsplit", ["1", "3"])
@pytest.mark.parametrize("variant", ["row_ring", "direct_sweep", "general"])
def test_backward_kernel_variants(monkeypatch, variant, gsplit):
    """The three backward kernels against the oracle: the row-ring sweep kernel (default for three-decision tables whose row
    windows slide upwards), the direct-load sweep kernel it falls back to, and the general kernel (any table width); each
    with one CTA sweeping the whole grid of its path tile and with the grid split over several CTAs."""
    if variant == "direct_sweep":
        monkeypatch.setenv("LSMC_SWEEP_RING", "0")
    elif variant == "general":
        monkeypatch.setenv("LSMC_SWEEP", "0")
    monkeypatch.setenv("LSMC_GSPLIT", gsplit)
    check_case(cases.three_factor_ratchets(num_sims=1500, num_days=40, seed=13))
    check_case(cases.simple_one_factor(num_sims=1009, num_days=30, seed=6))


def test_grid_split_of_direct_kernel(monkeypatch):
    """Direct-load backward kernel with one CTA sweeping all inventory-grid sub-tiles of its path tile (the layout used at
    production path counts) instead of one sub-tile per CTA."""
    monkeypatch.setenv("LSMC_GSPLIT", "1")
    check_case(cases.three_factor_ratchets(num_sims=1500, num_days=40, seed=14))


def test_caller_simulation_panels_larger_than_one_staging_chunk_round_trip():
    """UseSpotSimResults at a size where every panel spans several 64 MB chunks of the pinned staging pipeline
    (lsmc_set_sims in, lsmc_copy_panel out): what comes back is bit for bit what went in, and the valuation runs."""
    import time
    case = cases.simple_one_factor(num_sims=400_000, num_days=50, seed=3,
                                   flags=SimulationDataReturned.AllSpotPrices | SimulationDataReturned.Inventory)
    assert case.reg_spot.nbytes > 2 * (64 << 20)
    t0 = time.perf_counter()
    res = cases.run_gpu(case, keep_decisions=False)
    dt = time.perf_counter() - t0
    np.testing.assert_array_equal(res.RegressionSpotPriceSim, case.reg_spot)
    np.testing.assert_array_equal(res.ValuationSpotPriceSim, case.val_spot)
    assert res.InventoryBySim.shape == (case.flat.num_periods + 1, 400_000)
    assert np.isfinite(res.Npv)
    print(f"400k x 55 panels in and out, valuation included: {dt:.2f} s")
