"""Worker of tests/test_gpu_sharded.py: one process per GPU.  Every rank values its slice of ONE global set of pre-simulated
paths through the public call (LsmcStorageValuation.Calculate with a PathShard); rank 0 also values the whole set alone."""
import ctypes as C
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make_case(kind, total):
    from tests import cases
    if kind == "ratchets":
        return cases.three_factor_ratchets(num_sims=total, num_days=45, seed=41)
    if kind == "terminal":  # !MustBeEmptyAtEnd: terminal inventory value incl. the reference's running-sum quirk
        return cases.simple_one_factor(num_sims=total, num_days=30, seed=42, terminal=(-999.0, 0.0, 0.0, 1.0))
    if kind == "current_active":  # period 0 regresses on the sample mean (mean mode of the exchange)
        return cases.simple_one_factor(num_sims=total, num_days=30, seed=43, val_offset=7)
    raise ValueError(kind)


def summary(res, periods):
    trig = []
    for p in periods[:-1]:
        t = res.TriggerPriceVolumeProfiles[p]
        trig.append([x for pair in t.InjectTriggerPrices + t.WithdrawTriggerPrices for x in pair])
    return {"npv": res.Npv, "stderr": res.ValuationSimStandardError, "backward_npv": res.BackwardNpv,
            "deltas": np.array([res.Deltas[p] for p in periods]),
            "profile": np.array([[sp.Inventory, sp.InjectWithdrawVolume, sp.CmdtyConsumed, sp.InventoryLoss, sp.PeriodPv]
                                 for sp in (res.ExpectedStorageProfile[p] for p in periods)]),
            "trigger": trig, "pv_by_sim": res.PvBySim.copy(), "snapped": res.SnappedInventories}


def worker(rank, world, port, kind, total, out_dir, env):
    sys.path.insert(0, ROOT)
    os.environ.update(env)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    from storage_b200 import (LsmcStorageValuation, PathShard, SimulationDataReturned, SpotSimResultsFromPanels, _abi)
    lib = _abi.load_library()
    ident = torch.zeros(_abi.LSMC_COMM_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        buf = (C.c_ubyte * _abi.LSMC_COMM_ID_BYTES)()
        assert lib.lsmc_comm_unique_id(buf) == 0
        ident = torch.tensor(list(buf), dtype=torch.uint8)
    dist.broadcast(ident, 0)
    shard = PathShard(rank, world, bytes(ident.tolist()), total)

    case = make_case(kind, total)  # same seeds on every rank: identical global panels
    periods = case.flat.periods
    b = case.builder.Clone()
    sl = slice(shard.begin, shard.end)
    b.UseSpotSimResults(SpotSimResultsFromPanels(case.sim_periods, case.reg_spot[:, sl], [f[:, sl] for f in case.reg_factors]),
                        SpotSimResultsFromPanels(case.sim_periods, case.val_spot[:, sl], [f[:, sl] for f in case.val_factors]))
    b.SimulationDataReturned = SimulationDataReturned.Nothing
    valuation = LsmcStorageValuation(device=rank, shard=shard, reuse_handle=True)
    res = valuation.Calculate(b.Build())
    out = {"sharded": summary(res, periods), "begin": shard.begin, "end": shard.end}
    # regression coefficients must be bit-identical on every rank
    h = valuation._kept[1]
    P, Gmax, K = case.flat.num_periods, lib.lsmc_max_grid_points(h.h), case.flat.powers.shape[0]
    gamma = np.zeros((P, Gmax, K))
    h.check(lib.lsmc_get_regression(h.h, gamma.ctypes.data_as(_abi.c_double_p), None))
    out["gamma"] = gamma
    # second run on the kept handle (sequence numbers of the exchange keep increasing)
    res2 = valuation.Calculate(b.Build())
    out["second_npv"] = res2.Npv
    valuation.Close()
    # device simulator: a rank's draws are the corresponding slice of the unsharded draws (Philox keyed by global path id)
    bm = case.builder.Clone()
    bm.SimulationDataReturned = SimulationDataReturned.SpotPricesForRegression
    bm.SimulateWithMultiFactorModelAndMersenneTwister(case.model, total, 7, 8)
    sim = LsmcStorageValuation(device=rank, shard=shard).Calculate(bm.Build())
    out["sim_spot"] = sim.RegressionSpotPriceSim
    out["sim_npv"] = sim.Npv
    dist.barrier()
    if rank == 0:
        single = LsmcStorageValuation(device=0).Calculate(case.params)
        out["single"] = summary(single, periods)
        full = LsmcStorageValuation(device=0).Calculate(bm.Build())
        out["single_sim_spot"] = full.RegressionSpotPriceSim
        out["single_sim_npv"] = full.Npv
    with open(os.path.join(out_dir, f"rank{rank}.pkl"), "wb") as f:
        pickle.dump(out, f)
    dist.barrier()
    dist.destroy_process_group()
