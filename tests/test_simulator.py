"""Device multi-factor simulator vs the analytic moments of the model (north star: "agree with the reference simulator's
moments within stated Monte Carlo error").  The analytic covariance is MultiFactorModel.integrated_covar of the reference
(src/Cmdty.Storage.Python/cmdty_storage/multi_factor_diffusion_model.py:49-70):
  Cov(ln F(t,T1), ln F(t,T2)) = sum_ij vol_i(T1) vol_j(T2) rho_ij exp(-a_i T1 - a_j T2) (exp((a_i+a_j) t) - 1)/(a_i+a_j)."""
import ctypes as C
import math

import numpy as np
import pytest

from storage_b200 import _abi
from storage_b200.errors import check_status


def integrated_covar(mr, vol1, vol2, corr, t_obs, t1, t2):
    cov = 0.0
    for i in range(len(mr)):
        for j in range(len(mr)):
            a = mr[i] + mr[j]
            ext = t_obs if a == 0.0 else (math.exp(a * t_obs) - 1.0) / a
            cov += vol1[i] * vol2[j] * corr[i][j] * math.exp(-mr[i] * t1 - mr[j] * t2) * ext
    return cov


MR = np.array([16.2, 2.5, 0.0])
CORR = np.array([[1.0, 0.6, 0.3], [0.6, 1.0, 0.4], [0.3, 0.4, 1.0]])  # test_multi_factor_spot_sim.py:46-50


def model_arrays(P):
    t = (np.arange(P + 1) + 3.0) / 365.0
    vol = np.stack([np.linspace(0.95, 0.89, P + 1), np.linspace(0.15, 0.21, P + 1), np.linspace(0.35, 0.32, P + 1)])
    fwd = 56.85 + 5.0 * np.sin(np.arange(P + 1) / 20.0)
    return t, np.ascontiguousarray(vol), fwd


def simulate(P, N, seed, antithetic, first_path_id=0):
    import torch
    lib = _abi.load_library()
    t, vol, fwd = model_arrays(P)
    mp = _abi.lsmc_multi_factor_params()
    mp.num_factors = 3
    mp.mean_reversion = MR.ctypes.data_as(_abi.c_double_p)
    mp.factor_vol = vol.ctypes.data_as(_abi.c_double_p)
    corr = np.ascontiguousarray(CORR)
    mp.factor_corr = corr.ctypes.data_as(_abi.c_double_p)
    mp.period_time = t.ctypes.data_as(_abi.c_double_p)
    spot = torch.empty((P + 1, N), dtype=torch.float64, device="cuda")
    fac = [torch.empty((P + 1, N), dtype=torch.float64, device="cuda") for _ in range(3)]
    ptrs = (C.c_void_p * 3)(*[f.data_ptr() for f in fac])
    check_status(lib.lsmc_simulate_noalloc(0, C.byref(mp), P, 0, fwd.ctypes.data_as(_abi.c_double_p), N, seed, int(antithetic),
                                           first_path_id, spot.data_ptr(), ptrs, None), None)
    return spot.cpu().numpy(), np.stack([f.cpu().numpy() for f in fac]), (t, vol, fwd)


@pytest.mark.gpu
def test_moments_match_the_analytic_model():
    P, N = 40, 400_000
    spot, fac, (t, vol, fwd) = simulate(P, N, seed=12, antithetic=False)
    for r in (0, 7, 25, 40):
        s = spot[r]
        se = s.std() / math.sqrt(N)
        assert abs(s.mean() - fwd[r]) < 4.5 * se, f"E[S] != F at row {r}"  # martingale: E[S_T] = F(0,T)
        var = integrated_covar(MR, vol[:, r], vol[:, r], CORR, t[r], t[r], t[r])
        logs = np.log(s)
        # Var of a sample variance of a normal: 2 sigma^4/(N-1)
        assert abs(logs.var() - var) < 5 * var * math.sqrt(2.0 / N)
        assert abs(logs.mean() - (math.log(fwd[r]) - 0.5 * var)) < 4.5 * math.sqrt(var / N)
    # covariance of log-spot between two delivery dates = covariance of the factor states, mean-reverted
    r1, r2 = 10, 30
    x1, x2 = fac[:, r1, :], fac[:, r2, :]
    for i in range(3):
        for j in range(3):
            a = MR[i] + MR[j]
            expected = CORR[i, j] * math.exp(-MR[j] * (t[r2] - t[r1])) * ((1 - math.exp(-a * t[r1])) / a if a else t[r1])
            sample = np.mean(x1[i] * x2[j])
            sd = math.sqrt(np.var(x1[i] * x2[j]) / N)
            assert abs(sample - expected) < 5 * sd + 1e-12, (i, j, sample, expected)


@pytest.mark.gpu
def test_antithetic_pairs_and_sharding_independence():
    P, N = 12, 4096
    spot, fac, _ = simulate(P, N, seed=7, antithetic=True)
    np.testing.assert_array_equal(fac[:, :, 0::2], -fac[:, :, 1::2])  # paths 2j and 2j+1 use z and -z
    # a path's draws depend on its global id only: two shards reproduce the single run
    a, fa, _ = simulate(P, N // 2, seed=7, antithetic=True, first_path_id=0)
    b, fb, _ = simulate(P, N // 2, seed=7, antithetic=True, first_path_id=N // 2)
    np.testing.assert_array_equal(np.concatenate([a, b], axis=1), spot)
    np.testing.assert_array_equal(np.concatenate([fa, fb], axis=2), fac)
    c, _, _ = simulate(P, N, seed=8, antithetic=True)
    assert not np.array_equal(c, spot)


def test_host_generator_used_for_parity_inputs_has_the_model_moments():
    """The numpy generator that feeds identical paths to oracle and GPU (tests/cases.py) is itself checked against the
    analytic variance, so the parity inputs are model-consistent."""
    from tests import cases
    from storage_b200 import Day
    model = cases.MultiFactorParameters.For1Factor(12.5, 0.95)
    cur = Day(2019, 8, 29)
    periods = [cur + 5, cur + 60]
    fwd = {p: 50.0 for p in periods}
    spot, fac = cases.simulate_host(model, cur, periods, fwd, 200_000, seed=3)
    for r, p in enumerate(periods):
        tt = (p.ordinal - cur.ordinal) / 365.0
        var = 0.95 ** 2 * (1 - math.exp(-25.0 * tt)) / 25.0  # TestHelper.cs:76-85 one-factor variance
        assert abs(np.log(spot[r]).var() - var) < 6 * var * math.sqrt(2.0 / 200_000)
        assert abs(spot[r].mean() - 50.0) < 5 * spot[r].std() / math.sqrt(200_000)
