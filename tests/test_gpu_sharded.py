"""Sharded valuation on two GPUs, one process per GPU, through the C ABI: lsmc_comm_init, the sharded Gram
gather, the per-step exchange of the regression sums (peer-memory kernel and the NCCL all-gather fallback), the rank-ordered
sums of the forward statistics, the sharded terminal inventory value — against the single-GPU run of the same global paths.

Needs two GPUs: skipped on a one-GPU box (run with `gpurun --gpus 2`)."""
import os
import pickle
import socket
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(kind, total, env):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from tests.sharded_worker import worker
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(worker, args=(2, _free_port(), kind, total, d, env), nprocs=2, join=True)
        return [pickle.load(open(os.path.join(d, f"rank{r}.pkl"), "rb")) for r in range(2)]


def _check(outs, rel=1e-12):
    single = outs[0]["single"]
    for o in outs:
        s = o["sharded"]
        assert s["npv"] == pytest.approx(single["npv"], rel=rel)
        assert s["stderr"] == pytest.approx(single["stderr"], rel=1e-10)
        assert s["backward_npv"] == pytest.approx(single["backward_npv"], rel=rel)
        assert np.abs(s["deltas"] - single["deltas"]).max() <= rel * np.abs(single["deltas"]).max()
        for col in range(5):
            scale = np.abs(single["profile"][:, col]).max() + 1e-300
            assert np.abs(s["profile"][:, col] - single["profile"][:, col]).max() <= rel * scale
        for a, b in zip(s["trigger"], single["trigger"]):
            np.testing.assert_allclose(a, b, rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(s["pv_by_sim"], single["pv_by_sim"][o["begin"]:o["end"]], rtol=1e-10, atol=1e-7)
        assert o["second_npv"] == s["npv"]  # a second run on the same handle repeats the first to the bit
    # every rank holds the same results and the same regression coefficients, to the bit
    assert outs[0]["sharded"]["npv"] == outs[1]["sharded"]["npv"]
    np.testing.assert_array_equal(outs[0]["gamma"], outs[1]["gamma"])
    np.testing.assert_array_equal(outs[0]["sharded"]["deltas"], outs[1]["sharded"]["deltas"])
    assert outs[0]["sharded"]["snapped"] == outs[1]["sharded"]["snapped"] == single["snapped"]
    # the device simulator's draws do not depend on the sharding
    full = outs[0]["single_sim_spot"]
    for o in outs:
        np.testing.assert_array_equal(o["sim_spot"], full[:, o["begin"]:o["end"]])
        assert o["sim_npv"] == pytest.approx(outs[0]["single_sim_npv"], rel=rel)


@pytest.mark.parametrize("exchange", ["peer_memory", "nccl_all_gather"])
def test_two_ranks_equal_one_gpu_ratchets(exchange):
    _check(_run("ratchets", 3000, {"LSMC_PEER_EXCHANGE": "1" if exchange == "peer_memory" else "0"}))


def test_two_ranks_terminal_inventory_value():
    _check(_run("terminal", 2000, {}))


def test_two_ranks_current_period_active():
    _check(_run("current_active", 2000, {}))
