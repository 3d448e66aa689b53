"""GPU parity at the BASELINE.json configurations' own sizes (the large-N code paths — thousands of path tiles,
one CTA sweeping the whole grid, the [G][tiles][8] partial-sum reduction over thousands of chunks, pitch != N, the 64-chunk Gram
reduction — must meet the oracle too).

  configs[1]  50,000 paths x 365 daily steps, 100-point grid, 3-factor basis      exact + qr oracle
  configs[2]  200,000 paths x 365 daily steps, 100-point grid, ratchets/outages    qr oracle (the reference's algorithm)
  configs[3]  shape at 1,000,000 paths x 30 daily steps, 200-point grid            qr oracle

Same assertions as tests/test_gpu_parity.py (every argmax-index flip is a tie between equal volumes; no volume differs;
NPV / deltas / profile / per-sim panels within 1e-9; trigger prices).  The OpenMP oracle needs roughly 10-60 s per case on the
GPU box's host cores.
"""
import pytest

from tests import cases
from tests.test_gpu_parity import check_case

pytestmark = pytest.mark.gpu


def test_config1_full_size_50k_paths_365_steps():
    check_case(cases.three_factor_ratchets(num_sims=50_000, num_days=365, seed=31, outages=False, grid_points=100), modes=("exact", "qr"))


def test_config2_full_size_200k_paths_365_steps():
    check_case(cases.three_factor_ratchets(num_sims=200_000, num_days=365, seed=32, grid_points=100), modes=("qr",))


def test_config3_shape_1m_paths_30_steps_200_grid_points():
    check_case(cases.three_factor_ratchets(num_sims=1_000_000, num_days=30, seed=33, grid_points=200), modes=("qr",))
