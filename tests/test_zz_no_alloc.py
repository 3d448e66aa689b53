"""lsmc_simulate_noalloc — the simulator entry point for hosts that may not allocate GPU memory while running (services
that reserve their device memory once at start-up, hosts that manage the GPU heap through their own arena).  It writes into
buffers the caller provides and keeps its per-period parameter table in a fixed array inside the library instead of
cudaMalloc'ing one.

The test drives the new entry point through the C ABI twice in a row, with different horizons, seeds and path counts, so
the second call finds the first call's rows still in the shared table, and compares each result with what the existing
entry point (lsmc_simulate on a handle, read back with lsmc_copy_panel) produces for the same inputs.  Both run the same
arithmetic on the same draws, so the outputs must be identical to the last bit.  This file sorts after every other test
file on purpose: a run that stops at the first failure has then already run everything else.
"""
import ctypes as C

import numpy as np
import pytest

from storage_b200 import FlatProblem, Handle, _abi
from storage_b200.errors import check_status
from tests import cases

pytestmark = pytest.mark.gpu


def _existing_entry_point(case, seed, first_path_id):
    """Simulations from lsmc_simulate (allocating) for the case's model: spot [rows][N] and factors [F][rows][N]."""
    lib = _abi.load_library()
    flat = FlatProblem(case.params, case.num_sims, case.model.NumFactors)
    h = Handle(flat, 0)
    try:
        mp, keep = flat.model_arrays(case.model, case.params.CurrentPeriod)
        h.check(lib.lsmc_simulate(h.h, 0, C.byref(mp), seed, 1, first_path_id))
        h.check(lib.lsmc_simulate(h.h, 1, C.byref(mp), seed + 1, 1, first_path_id))
        h.check(lib.lsmc_run(h.h, None, None, None))  # panels can be read back once a run has completed
        rows = lib.lsmc_num_sim_periods(h.h)
        spot = h.panel(_abi.LSMC_PANEL_SPOT_REGRESSION, rows, case.num_sims)
        factors = h.panel(_abi.LSMC_PANEL_FACTORS_REGRESSION, rows, case.num_sims, case.model.NumFactors)
        del keep
        return flat, spot, factors
    finally:
        h.close()


def _no_alloc_entry_point(case, flat, seed, first_path_id):
    import torch
    lib = _abi.load_library()
    F, N = case.model.NumFactors, case.num_sims
    rows = flat.num_periods + 1 - flat.first_sim
    # the caller owns every device buffer; torch stands in for the host's own arena here
    spot = torch.full((rows, N), float("nan"), dtype=torch.float64, device="cuda")
    factors = [torch.full((rows, N), float("nan"), dtype=torch.float64, device="cuda") for _ in range(F)]
    ptrs = (C.c_void_p * F)(*[f.data_ptr() for f in factors])
    mp, keep = flat.model_arrays(case.model, case.params.CurrentPeriod)
    torch.cuda.synchronize()
    status = lib.lsmc_simulate_noalloc(0, C.byref(mp), flat.num_periods, flat.first_sim, flat.forward.ctypes.data_as(_abi.c_double_p),
                                       N, seed, 1, first_path_id, spot.data_ptr(), ptrs, None)
    check_status(status, None)
    del keep
    return spot.cpu().numpy(), np.stack([f.cpu().numpy() for f in factors])


def test_two_calls_in_a_row_match_the_allocating_entry_point():
    first = cases.three_factor_ratchets(num_sims=4096, num_days=45, seed=1, outages=False)
    second = cases.simple_one_factor(num_sims=1536, num_days=20, seed=2, val_offset=7)  # shorter horizon, 1 factor, active period
    for case, seed, first_path in ((first, 1234, 0), (second, 99, 4096)):
        flat, spot_ref, factors_ref = _existing_entry_point(case, seed, first_path)
        spot, factors = _no_alloc_entry_point(case, flat, seed, first_path)
        assert np.isfinite(spot).all() and np.isfinite(factors).all()
        np.testing.assert_array_equal(spot, spot_ref)
        np.testing.assert_array_equal(factors, factors_ref)


def test_horizon_beyond_the_fixed_table_is_rejected():
    import torch
    from storage_b200 import ArgumentException
    lib = _abi.load_library()
    limit = lib.lsmc_simulate_noalloc_max_periods()
    assert limit == 2048
    P = limit + 5
    F = 1
    mr = np.array([1.0])
    vol = np.full((F, P + 1), 0.3)
    corr = np.eye(F)
    t = np.arange(P + 1, dtype=np.float64) / 365.0
    fwd = np.full(P + 1, 50.0)
    mp = _abi.lsmc_multi_factor_params()
    mp.num_factors = F
    mp.mean_reversion = mr.ctypes.data_as(_abi.c_double_p)
    mp.factor_vol = vol.ctypes.data_as(_abi.c_double_p)
    mp.factor_corr = corr.ctypes.data_as(_abi.c_double_p)
    mp.period_time = t.ctypes.data_as(_abi.c_double_p)
    out = torch.zeros(8, dtype=torch.float64, device="cuda")
    ptrs = (C.c_void_p * 1)(out.data_ptr())
    status = lib.lsmc_simulate_noalloc(0, C.byref(mp), P, 0, fwd.ctypes.data_as(_abi.c_double_p), 4, 1, 0, 0, out.data_ptr(), ptrs, None)
    assert status == _abi.LSMC_ERR_ARGUMENT
    with pytest.raises(ArgumentException, match="exceed the fixed scratch"):
        check_status(status, None)
