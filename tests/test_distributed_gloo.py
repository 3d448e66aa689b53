"""World-size-2 checks of the path-sharding logic on CPU (gloo): shard bounds, rank-ordered merging of regression sums and
of the NPV / standard-error statistics, against a single-process oracle run.  The GPU library performs the same merges
with NCCL all-gathers (storage_b200/csrc/comm.cu); this covers the host-side logic without a GPU."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from storage_b200 import sharding
    from tests import cases
    from oracle import oracle
    case = cases.simple_one_factor(num_sims=total, num_days=15, seed=21)
    full = oracle.lsmc_calculate(case.flat, case.reg_spot, case.reg_factors, case.val_spot, case.val_factors, mode=0)
    begin, end = sharding.shard_bounds(total, world, rank)
    pv = torch.tensor(full.pv_by_sim[begin:end])
    # rank-ordered merge of the sums, as the library does after its all-gather
    sums = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(sums, pv.sum().reshape(1))
    npv = sharding.merge_npv([float(s) for s in sums], total)
    sq = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(sq, ((pv - npv) ** 2).sum().reshape(1))
    se = sharding.merge_standard_error([float(s) for s in sq], total)
    # regression sums: X^T y over the shard, merged in rank order, must not depend on which rank merges
    x = oracle.populate_design_matrix(case.flat.powers, case.reg_spot[3], [f[3] for f in case.reg_factors])
    local = torch.tensor(x[:, begin:end] @ full.pv_by_sim[begin:end])
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    merged = sharding.merge_regression_sums([g.tolist() for g in gathered])
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.array([npv, se, full.npv, full.std_error, begin, end] + merged))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_merge_matches_single_process(tmp_path):
    world, total = 2, 1001  # odd: the last pair is incomplete
    port = _free_port()
    mp.spawn(_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
    r0 = np.load(tmp_path / "rank0.npy")
    r1 = np.load(tmp_path / "rank1.npy")
    assert r0[4] == 0 and r1[5] == total and r0[5] == r1[4] and r0[5] % 2 == 0  # contiguous, pair-aligned shards
    assert r0[0] == pytest.approx(r0[2], rel=1e-13) and r0[1] == pytest.approx(r0[3], rel=1e-10)
    np.testing.assert_array_equal(r0[6:], r1[6:])  # every rank holds bit-identical merged regression sums
    assert r0[0] == r1[0] and r0[1] == r1[1]


def test_shard_bounds_cover_all_paths_once():
    from storage_b200.sharding import shard_bounds
    for total in (1, 2, 7, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            edges = [shard_bounds(total, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            assert all(b % 2 == 0 or b == total for b, _ in edges)
