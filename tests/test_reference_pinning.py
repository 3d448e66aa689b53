"""Pins the oracle — and with it everything that is checked against the oracle — on numbers the REFERENCE produced.

1. The reference's simulator golden (src/Cmdty.Storage.Python/tests/test_multi_factor_spot_sim.py:64-89: four simulations x
   three periods of a three-factor model, seed 12, seventeen-digit values) is reproduced by oracle/reference_sim.py, the
   restatement of the un-vendored chain MathNet MersenneTwister -> MathNet Normal.Samples -> Cmdty.Core
   MultiFactorSpotPriceSimulator.
2. Fed with the paths that chain generates for the reference's LSMC regression tests (test_multi_factor.py:34-124 and :199-270;
   500 antithetic paths, seed 11), the oracle's Calculate returns the reference's literal NPV and standard error TO THE LAST
   DIGIT, in both regression modes (`qr`, the reference's algorithm in double, and `exact`), and its deltas, expected profile
   and trigger prices equal the reference's CSV fixtures (copied verbatim to tests/golden/).

No GPU needed; tests/test_gpu_reference_goldens.py runs the same two cases through the CUDA library."""
import numpy as np
import pytest

from oracle import oracle, reference_sim
from tests import reference_goldens


def test_simulator_reproduces_the_reference_golden_values():  # test_multi_factor_spot_sim.py:33-89
    mean_reversion = [0.0, 2.5, 16.2]
    vol = np.array([[0.35, 0.29, 0.32], [0.15, 0.18, 0.21], [0.95, 0.92, 0.89]])  # [factor][period]
    corr = [[1.0, 0.6, 0.3], [0.6, 1.0, 0.4], [0.3, 0.4, 1.0]]
    forward = [56.85, 59.08, 62.453]
    times = [5 / 365.0, 172 / 365.0, 368 / 365.0]  # 2020-07-27 -> 2020-08-01, 2021-01-15, 2021-07-30, Act/365
    spot, _ = reference_sim.simulate(mean_reversion, vol, corr, times, forward, 4, 12, antithetic=False)
    golden = np.array([[52.59976397688973, 57.559631642935514, 89.40526992772634],
                       [46.1206448628463, 72.0381089486175, 85.18869803117379],
                       [58.15838580682589, 82.49607173562342, 138.68587285875978],
                       [65.500441945042979, 42.812676607997183, 76.586790647813046]])
    # the reference asserts exact equality; libm's exp/log may differ from .NET's in the last bit
    np.testing.assert_allclose(spot.T, golden, rtol=4e-16, atol=0)


@pytest.mark.parametrize("name", ["multi_factor_test-1", "three_factor_test-1"])
@pytest.mark.parametrize("mode", ["qr", "exact"])
def test_oracle_calculate_equals_the_reference_goldens(name, mode):
    case = reference_goldens.build(name)
    r = oracle.lsmc_calculate(case.flat, case.spot, case.factors, case.spot, case.factors, mode=0 if mode == "qr" else 1)
    # assertAlmostEqual(npv, golden, places=6) and assertEqual(golden, val_sim_standard_error) in the reference
    assert round(abs(r.npv - case.golden["npv"]), 6) == 0
    assert r.npv == pytest.approx(case.golden["npv"], rel=1e-15)
    assert r.std_error == pytest.approx(case.golden["std_error"], rel=1e-14)
    P = case.flat.num_periods
    trigger = [(t.has_inject_price, t.max_inject_volume, t.max_inject_trigger_price, t.has_withdraw_price, t.max_withdraw_volume,
                t.max_withdraw_trigger_price) for t in r.trigger[:P]]
    worst = reference_goldens.compare_with_csvs(name, r.deltas, r.profile, trigger, rel=1e-11)
    print(f"{name} [{mode}]: npv {r.npv!r} (golden {case.golden['npv']!r}), std error {r.std_error!r}; worst relative differences vs CSVs {worst}")
