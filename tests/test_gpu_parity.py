"""GPU parity: the CUDA path (through the C ABI / host API mirror) against the CPU oracle on identical pre-simulated paths
fed through the caller-supplied-simulation overload (north star).

Tolerances (BASELINE.json north_star): NPV, deltas, profile within 1e-9 relative in FP64; argmax decisions identical.
"Identical decisions" is checked on decision *volumes*: when the inventory-space bound makes two entries of a decision set
coincide to ~1e-12 (e.g. max-rate withdrawal vs withdrawal constrained by the next step's maximum inventory) the index of
the maximum is decided by the last ulp of the continuation value.  Every test therefore asserts: (1) wherever the argmax
INDEX differs from the oracle's, the two decisions' volumes coincide to 1e-9 of the largest rate (the flip is a tie), and
(2) no decision anywhere differs in volume by more than that.  The raw flip count of every case, against both oracle modes,
is printed (`PARITY {...}` lines), appended to gpurun_out/parity_report.jsonl and recorded in BASELINE.md §4.
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle
from storage_b200 import SimulationDataReturned
from tests import cases

pytestmark = pytest.mark.gpu

REL = 1e-9
VOLUME_TIE_REL = 1e-9  # two decisions are "the same volume" when they differ by less than this fraction of the largest rate


def decision_differences(gpu, ref):
    """(raw index flips, largest |volume difference| at a flipped position, largest anywhere, positions whose volumes differ
    by more than 1e-9 of the largest rate).  A raw flip is a different argmax INDEX at the same (period, path); it is a tie
    when the two decisions' volumes coincide to rounding (see the module docstring)."""
    idx_diff = gpu.DecisionIndexBySim != ref.decision_index
    dvol = np.abs(gpu.InjectWithdrawVolumeBySim[:-1] - ref.inject_withdraw_volume[:-1])
    scale = max(float(np.abs(ref.inject_withdraw_volume).max()), 1.0)
    raw = int(np.count_nonzero(idx_diff))
    at_flips = float(dvol[idx_diff].max()) if raw else 0.0
    material = int(np.count_nonzero(dvol > VOLUME_TIE_REL * scale))
    return raw, at_flips, float(dvol.max()), material, scale


_report_rows = []


def report(case_name, case, against, form, raw, at_flips, anywhere, material, npv_rel, delta_rel, snapped):
    """One line per (case, oracle mode, arithmetic form) on stdout (pytest -s / failure output) and in gpurun_out/."""
    row = {"case": case_name, "paths": int(case.num_sims), "periods": int(case.flat.num_periods), "oracle": against, "kernel_form": form,
           "decisions": int(case.num_sims) * int(case.flat.num_periods), "raw_index_flips": raw, "max_dvolume_at_flips": at_flips,
           "max_dvolume": anywhere, "material_differences": material, "npv_rel": npv_rel, "delta_rel": delta_rel,
           "inventories_snapped_onto_a_limit": snapped}
    _report_rows.append(row)
    print("PARITY " + json.dumps(row))
    try:
        out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
        os.makedirs(out_dir, exist_ok=True)
        with open(os.path.join(out_dir, "parity_report.jsonl"), "a") as f:
            f.write(json.dumps(row) + "\n")
    except OSError:
        pass


def check_case(case, rel=REL, modes=("exact", "qr"), name=None):
    """Both arithmetic forms of the backward kernel (default fused payoff, and the reference's operation order) against both
    oracle modes: `exact` (the least-squares fit evaluated in double-double, rounded once) and `qr` (the reference's own
    algorithm, thin QR -> R^-1 -> pinv in double, LSV.cs:193-206)."""
    import inspect
    name = name or inspect.stack()[1].function
    case.params.SimulationDataReturned = SimulationDataReturned.All
    refs = {m: oracle.lsmc_calculate(case.flat, case.reg_spot, case.reg_factors, case.val_spot, case.val_factors,
                                     mode=1 if m == "exact" else 0) for m in modes}
    for reference_order in (False, True):
        gpu = cases.run_gpu(case, reference_operation_order=reference_order)
        for m in modes:
            compare_with_oracle(case, gpu, refs[m], rel, name=name, against=m, form="reference-order" if reference_order else "fused")
    return gpu, refs[modes[0]]


def compare_with_oracle(case, gpu, ref, rel, name="", against="exact", form="fused"):
    raw, at_flips, anywhere, material, scale = decision_differences(gpu, ref)
    periods = case.flat.periods
    deltas = np.array([gpu.Deltas[p] for p in periods])
    dscale = np.abs(ref.deltas).max() + 1e-300
    report(name, case, against, form, raw, at_flips, anywhere, material, abs(gpu.Npv - ref.npv) / max(abs(ref.npv), 1e-300),
           float(np.abs(deltas - ref.deltas).max() / dscale), int(gpu.SnappedInventories))
    # the documented departure (inventory moved onto a limit where the reference throws) is taken equally often on both sides;
    # a tie flip changes the inventory by rounding and with it, possibly, whether the next period needs the move
    assert abs(gpu.SnappedInventories - ref.snapped_inventories) <= raw, (gpu.SnappedInventories, ref.snapped_inventories)
    # argmax indices: every flip must be a tie (the two decisions inject/withdraw the same volume to rounding), and no
    # position anywhere may differ in volume
    assert at_flips <= VOLUME_TIE_REL * scale, f"{raw} index flips, one between volumes {at_flips:.3e} apart (not a tie)"
    assert material == 0, f"{material} decisions differ in volume by more than {VOLUME_TIE_REL:g} of the rate ({raw} raw index flips)"
    assert gpu.Npv == pytest.approx(ref.npv, rel=rel)
    assert gpu.ValuationSimStandardError == pytest.approx(ref.std_error, rel=1e-8)
    assert gpu.BackwardNpv == pytest.approx(ref.backward_npv, rel=rel)
    P = case.flat.num_periods
    assert np.abs(deltas - ref.deltas).max() <= rel * dscale
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


@pytest.mark.parametrize("variant", ["row_ring", "row_ring_dfma", "direct_sweep", "general"])
@pytest.mark.parametrize("num_basis", [10, 13])
def test_wide_bases(monkeypatch, variant, num_basis):
    """Bases of 10 and 13 monomials: the kernels' 12- and 16-wide instantiations (two 8-row blocks of regression sums per grid
    block, 16-wide partial-sum rows), in all three backward kernels."""
    from storage_b200 import BasisFunctionsBuilder, Sim
    if variant == "row_ring_dfma":
        monkeypatch.setenv("LSMC_SWEEP_MMA", "0")
    elif variant == "direct_sweep":
        monkeypatch.setenv("LSMC_SWEEP_RING", "0")
    elif variant == "general":
        monkeypatch.setenv("LSMC_SWEEP", "0")
    basis = (BasisFunctionsBuilder.Ones + Sim.X0 + Sim.X0.Pow(2) + Sim.X1 + Sim.X2 + Sim.X0 * Sim.X1 + Sim.S + Sim.S.Pow(2)
             + Sim.X1.Pow(2) + Sim.X2.Pow(2))
    if num_basis == 13:
        basis = basis + Sim.X0 * Sim.X2 + Sim.X1 * Sim.X2 + Sim.S * Sim.X0
    check_case(cases.three_factor_ratchets(num_sims=3000, num_days=30, seed=17, basis=basis))


@pytest.mark.parametrize("gsplit", ["1", "3"])
@pytest.mark.parametrize("variant", ["row_ring", "row_ring_dfma", "direct_sweep", "general"])
def test_backward_kernel_variants(monkeypatch, variant, gsplit):
    """The backward kernels against the oracle: the row-ring sweep kernels (default for three-decision tables whose row windows
    slide upwards: FP64-mma evaluation for the fused form, DFMA evaluation for the reference's operation order and, with
    LSMC_SWEEP_MMA=0, for both), the direct-load sweep kernel they fall back to, and the general kernel (any table width);
    each with one CTA sweeping the whole grid of its path tile and with the grid split over several CTAs."""
    if variant == "row_ring_dfma":
        monkeypatch.setenv("LSMC_SWEEP_MMA", "0")
    elif variant == "direct_sweep":
        monkeypatch.setenv("LSMC_SWEEP_RING", "0")
    elif variant == "general":
        monkeypatch.setenv("LSMC_SWEEP", "0")
    monkeypatch.setenv("LSMC_GSPLIT", gsplit)
    if variant == "row_ring":  # the FP64-mma kernel: whole tiles only / some whole tiles, the others in three pieces
        monkeypatch.setenv("LSMC_MMA_TAIL_SPLIT", "1000000,1" if gsplit == "1" else "5,3")
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


@pytest.mark.parametrize("block", ["7", "64"])
def test_streamed_panels_equal_resident_panels(monkeypatch, block):
    """LSMC_STREAM_PANELS: the forward simulation run in blocks of periods, every block's panel rows copied to the
    caller's arrays while the next block is computed, the per-path state carried between the launches — bit for bit what the
    one-launch forward simulation with device-resident panels returns (all six result panels, the decision indices, and the
    scalar results).  Seven-period blocks exercise many ring turns and a ragged last block."""
    monkeypatch.setenv("LSMC_STREAM_BLOCK", block)
    for case in (cases.three_factor_ratchets(num_sims=1500, num_days=40, seed=51, flags=SimulationDataReturned.All),
                 cases.simple_one_factor(num_sims=1009, num_days=30, seed=52, flags=SimulationDataReturned.All, terminal=(-999.0, 0.0, 0.0, 1.0))):
        from storage_b200 import LsmcStorageValuation
        resident = LsmcStorageValuation().Calculate(case.params, keep_decisions=True, stream_panels=False)
        streamed = LsmcStorageValuation().Calculate(case.params, keep_decisions=True, stream_panels=True)
        assert streamed.Npv == resident.Npv and streamed.ValuationSimStandardError == resident.ValuationSimStandardError
        for name in ("InventoryBySim", "InjectWithdrawVolumeBySim", "CmdtyConsumedBySim", "InventoryLossBySim", "NetVolumeBySim",
                     "PvByPeriodAndSim", "DecisionIndexBySim", "PvBySim"):
            np.testing.assert_array_equal(getattr(streamed, name), getattr(resident, name), err_msg=name)
        assert list(streamed.Deltas.values()) == list(resident.Deltas.values())


def test_forward_table_staging_modes_agree(monkeypatch):
    """The forward kernel stages a period's tables in one shared-memory buffer (many path tiles per CTA) or in two, the next
    period's tables fetched under the current period's arithmetic (one or two tiles per CTA): same arithmetic, so bit-identical
    results, resident and with streamed panels (the launch of a period block starts a fresh buffer sequence)."""
    from storage_b200 import LsmcStorageValuation
    monkeypatch.setenv("LSMC_STREAM_BLOCK", "7")
    for case in (cases.three_factor_ratchets(num_sims=1500, num_days=40, seed=61, flags=SimulationDataReturned.All),
                 cases.simple_one_factor(num_sims=1009, num_days=30, seed=62, flags=SimulationDataReturned.All)):
        results = {}
        for buffers in ("1", "2"):
            monkeypatch.setenv("LSMC_FWD_TABLE_BUFFERS", buffers)
            results[buffers] = [LsmcStorageValuation().Calculate(case.params, keep_decisions=True, stream_panels=s) for s in (False, True)]
        one = results["1"][0]
        for res in (results["1"][1], results["2"][0], results["2"][1]):
            assert res.Npv == one.Npv and res.ValuationSimStandardError == one.ValuationSimStandardError
            for name in ("InventoryBySim", "InjectWithdrawVolumeBySim", "PvByPeriodAndSim", "DecisionIndexBySim", "PvBySim"):
                np.testing.assert_array_equal(getattr(res, name), getattr(one, name), err_msg=name)
            assert list(res.Deltas.values()) == list(one.Deltas.values())
        check_case(case)  # and the default mode against the oracle
