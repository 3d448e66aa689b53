#!/usr/bin/env python
"""bench.py — LSMC path*step*inventory-point evaluations per second on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W [--config C]     our arm (CUDA library through the C ABI)
  python bench.py --impl reference --steps K --warmup W           the reference algorithm's CPU port (oracle) on the host cores

A "step" is one full pass of the hot path — backward induction + forward simulation (lsmc_run) — over one batch of simulated
paths already resident in HBM.  --config picks one of BASELINE.json's shapes; the default, 4, is the one the
metric is quoted on: 5-year daily, 3-factor seasonal model, 1M paths, 200-point inventory grid, ratchets with outages, loss,
consumed commodity, inventory cost, empty at end.

Scaling is STRONG: the configuration's paths are a fixed total, sharded over the N ranks (one process per GPU; per backward
step the ranks exchange the G' x K regression sums).  With N > 1 the line also carries `weak` (every rank holding the whole
single-GPU path count, the figure round 1 reported) and `sharded_parity` (a small fixed-total case valued by the N ranks and
again by rank 0 alone on the same global path ids).

`e2e` is the same metric through the public call, LsmcStorageValuation.Calculate (storage_b200/lsmc.py, the mirror of
LsmcStorageValuation.cs:57), from host descriptors to host results: device simulation of both path sets, the run, and the
device->host copy of the results, all timed; `e2e_panels` is the caller-supplied-simulation overload (UseSpotSimResults) with
the panels in host memory, host->device copy timed.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

HBM_FALLBACK_GBS = 6650.0  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent
UNIT = "path*step*inventory-point evals/s"
METRIC = "LSMC path*step*inventory-point evals/sec"

# BASELINE.json configs[0..4] as SURVEY.md 8d tabulates them (cfg 1..5)
CONFIGS = {
    1: dict(sims=2_000, days=365, grid=100, kind="one_factor", outages=False, factors=1, basis=4,
            what="configs[0]: constant-parameter storage, 1-factor mean-reverting model, basis 1+s+s^2+s^3"),
    2: dict(sims=50_000, days=365, grid=100, kind="three_factor", outages=False, factors=3, basis=7,
            what="configs[1]: 3-factor seasonal model, basis 1+x0+x0^2+x1+x2+x0*x1+s, inventory-varying ratchets"),
    3: dict(sims=200_000, days=365, grid=100, kind="three_factor", outages=True, factors=3, basis=7,
            what="configs[2]: time/inventory-varying ratchets with outages, loss, consumed cmdty, inventory cost, empty at end"),
    4: dict(sims=1_000_000, days=1825, grid=200, kind="three_factor", outages=True, factors=3, basis=7,
            what="configs[3]: 5-year daily 3-factor valuation, storage as configs[2]"),
    5: dict(sims=4_000_000, days=730, grid=100, kind="three_factor", outages=False, factors=3, basis=7,
            what="configs[4]: 2-year daily 3-factor scaling sweep, storage as configs[1]"),
}


def build_params(cfg, num_sims, seed=11, flags=None):
    """LsmcValuationParameters of a configuration, built through the host API mirror (built-in simulation overload)."""
    from storage_b200 import (BasisFunctionsBuilder, Day, FixedSpacingStateSpaceGridCalc, LsmcValuationParameters,
                              MultiFactorParameters, SimulationDataReturned, StorageHelper, settle_on_20th_of_following_month)
    from tests import cases
    start = Day(2020, 4, 1)
    end = start + cfg["days"]
    val_date = start - 1
    if cfg["kind"] == "one_factor":
        storage = cases.simple_storage(start, end)
        model = MultiFactorParameters.For1Factor(12.5, 0.95)
        basis = cases.ONE_FACTOR_BASIS
        inventory = 5_250.0
    else:
        storage = cases.ratchet_storage(start, end, outages=cfg["outages"], kind="linear")
        model = cases.three_factor_model(start, end)
        basis = BasisFunctionsBuilder.Parse(cases.THREE_FACTOR_BASIS_TEXT, 3)
        inventory = 5_250.0  # 10 % of the maximum inventory
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = inventory
    b.ForwardCurve = cases.forward_curve(val_date, end)
    b.Storage = storage
    b.SettleDateRule = settle_on_20th_of_following_month
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.055)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, cfg["grid"])
    b.BasisFunctions = basis
    b.SimulationDataReturned = SimulationDataReturned.Nothing if flags is None else flags
    b.SimulateWithMultiFactorModelAndMersenneTwister(model, num_sims, seed, seed * 2)
    return b, model


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


def measured_traffic(cfg_id, paths_per_gpu, grid):
    """DRAM bytes of one launch of the dominant kernel from the committed ncu capture (profiles/sweep_traffic.json), when that
    capture was taken at this launch shape; None otherwise (a scaled constant is not a measurement)."""
    try:
        with open(os.path.join(ROOT, "profiles", "sweep_traffic.json")) as f:
            t = json.load(f)
        if t["paths"] == paths_per_gpu and t["grid_points"] == grid:
            return float(t["dram_bytes_per_launch"]), t["source"]
    except Exception:
        pass
    return None, None


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def oracle_evals(flat, n):
    """path*step*grid-point evaluations of one valuation, counted on the oracle's own grids (the reference arm must not map
    the product library)."""
    from oracle import oracle
    total = 0
    cnt = C.c_int()
    for p in range(flat.num_periods):
        rc = oracle.lib().oracle_period_grid(C.byref(flat.struct), p, None, 0, C.byref(cnt))
        if rc != 0:
            raise RuntimeError(oracle.lib().oracle_last_error().decode())
        total += cnt.value
    return total * n


def cpu_baseline(cfg, cpu_sims):
    """The oracle (port of the reference algorithm, QR regression in double, OpenMP over grid points / paths) on a bounded
    sample of the same workload shape, on all the host cores this process may use."""
    from oracle import oracle
    from storage_b200 import FlatProblem
    from tests import cases
    n = cpu_sims
    b, model = build_params(cfg, n)
    params = b.Build()
    flat = FlatProblem(params, n, cfg["factors"])
    fwd = params.ForwardCurve
    reg_spot, reg_fac = cases.simulate_host(model, params.CurrentPeriod, flat.sim_periods, fwd, n, 11)
    val_spot, val_fac = cases.simulate_host(model, params.CurrentPeriod, flat.sim_periods, fwd, n, 23)
    evals = oracle_evals(flat, n)
    threads = host_threads()  # explicit: torchrun exports OMP_NUM_THREADS=1
    t0 = time.perf_counter()
    res = oracle.lsmc_calculate(flat, reg_spot, list(reg_fac), val_spot, list(val_fac), mode=0, threads=threads, panels=False)
    dt = time.perf_counter() - t0
    return {"value": evals / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{n} paths x {flat.num_periods} daily steps x {cfg['grid']}-point grid, same storage/model/basis as the "
                      f"GPU workload, QR regression in double, {dt:.1f} s",
            "seconds": dt, "npv": res.npv}, evals, dt


def workload_config(cfg_id, cfg, total_paths, world, cpu=False):
    return {"workload": "BASELINE.json %s; %d daily steps, %d-point inventory grid" % (cfg["what"], cfg["days"], cfg["grid"]),
            "config_id": cfg_id, "total_paths": total_paths, "paths_per_gpu": total_paths // max(world, 1),
            "steps_per_valuation": cfg["days"], "grid_points": cfg["grid"], "factors": cfg["factors"], "basis_functions": cfg["basis"],
            "parallelism": ("cpu-openmp" if cpu else f"paths sharded x{world} (fixed total), per-step exchange of the regression sums"),
            "cache": "inputs larger than L2: simulated paths and per-step value slabs exceed the 126 MB L2 at this size"
                     if total_paths // max(world, 1) * cfg["grid"] * 16 > 126e6 else
                     "per-step value slabs fit in the 126 MB L2 at this size (small configuration; the simulated paths stream from HBM)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU reference runs on rank 0 only
    cfg = CONFIGS[args.config]
    times = []
    for i in range(args.warmup + args.steps):
        base, evals, dt = cpu_baseline(cfg, args.cpu_sims)
        if i >= args.warmup:
            times.append(dt)
    ms = 1000.0 * sum(times) / len(times)
    value = evals / (ms / 1000.0)
    base["value"] = value
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 0, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": workload_config(args.config, cfg, args.cpu_sims, 1, cpu=True), "cpu_baseline": base,
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def run_gpu(args):
    import torch
    from storage_b200 import FlatProblem, Handle, LsmcStorageValuation, PathShard, _abi
    from storage_b200.sharding import shard_bounds
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: storage_b200 has no CPU fallback")
    # rank 0 prints exactly ONE line on stdout: anything a library writes there meanwhile (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    cfg = dict(CONFIGS[args.config])
    if args.sims:
        cfg["sims"] = args.sims
    if args.days:
        cfg["days"] = args.days
    if args.grid:
        cfg["grid"] = args.grid
    total = cfg["sims"]
    lib = _abi.load_library()

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    comm_id = {}

    def communicator_id():
        """Rank 0 draws a communicator id once, torch.distributed carries it to the other ranks; every handle then joins with the
        same id (the library creates its communicator on first use and shares it afterwards)."""
        if "raw" not in comm_id:
            ident = torch.zeros(_abi.LSMC_COMM_ID_BYTES, dtype=torch.uint8)
            if rank == 0:
                buf = (C.c_ubyte * _abi.LSMC_COMM_ID_BYTES)()
                if lib.lsmc_comm_unique_id(buf) != 0:
                    raise SystemExit("lsmc_comm_unique_id failed: " + lib.lsmc_last_error(None).decode())
                ident = torch.tensor(list(buf), dtype=torch.uint8)
            ident = ident.cuda()
            dist.broadcast(ident, 0)
            comm_id["raw"] = bytes(ident.cpu().tolist())
        return comm_id["raw"]

    def open_handle(cfg_, total_paths, sharded=True):
        """Descriptor flattening, table build, device buffers, (joining the ranks,) device simulation of this rank's paths."""
        w, r = (world, rank) if sharded else (1, 0)
        begin, end = shard_bounds(total_paths, w, r)
        n_local = end - begin
        b, model = build_params(cfg_, n_local)
        params = b.Build()
        flat_ = FlatProblem(params, n_local, cfg_["factors"])
        handle = Handle(flat_, local_rank)
        if w > 1:
            handle.check(lib.lsmc_comm_init(handle.h, communicator_id(), r, w, total_paths))
        mp, keep = flat_.model_arrays(model, params.CurrentPeriod)
        handle.check(lib.lsmc_simulate(handle.h, 0, C.byref(mp), 11, 1, begin))
        handle.check(lib.lsmc_simulate(handle.h, 1, C.byref(mp), 22, 1, begin))
        del keep
        return flat_, handle, n_local

    def one_run(handle, arrays=False):
        handle.check(lib.lsmc_run(handle.h, None, None, None))
        r = _abi.lsmc_results()
        keep = None
        if arrays:
            P_ = handle.flat.num_periods
            keep = (np.zeros(P_ + 1), np.zeros((P_ + 1, 5)))
            r.deltas = keep[0].ctypes.data_as(_abi.c_double_p)
            r.profile = keep[1].ctypes.data_as(_abi.c_double_p)
        handle.check(lib.lsmc_get_results(handle.h, C.byref(r)))
        return r, keep

    def timed_runs(handle, steps, warmup, sampler=None):
        for _ in range(warmup):
            one_run(handle)
        barrier()
        if sampler is not None:
            sampler.start()
        times, acc = [], dict(kernel_ms=0.0, launches=0, bk_launches=0, device_ms=0.0)
        r = None
        for _ in range(steps):
            barrier()
            t0 = time.perf_counter()
            r, _ = one_run(handle)
            barrier()
            times.append(time.perf_counter() - t0)
            acc["kernel_ms"] += r.ms_backward_kernel
            acc["bk_launches"] += r.backward_kernel_launches
            acc["launches"] += r.kernel_launches
            acc["device_ms"] += r.ms_device
        step_s = max_over_ranks(sum(times) / len(times))
        return step_s, r, acc

    # ---- resident-input arm (strong scaling: the configuration's paths, sharded)
    t_setup = time.perf_counter()
    flat, h, n_local = open_handle(cfg, total)
    setup_s = time.perf_counter() - t_setup
    evals_local = lib.lsmc_num_evaluations(h.h)
    evals_total = evals_local // n_local * total
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t_all = time.perf_counter()
    step_s, r, acc = timed_runs(h, args.steps, args.warmup, sampler)
    total_s = time.perf_counter() - t_all
    clocks = sampler.stop() if rank == 0 else None
    phases = {"gram_cholesky_ms": r.ms_pseudo_inverse, "backward_ms": r.ms_backward, "forward_ms": r.ms_forward, "run_total_ms": r.ms_total}
    npv = r.npv
    P = flat.num_periods
    h.close()

    # ---- end to end through the public call, LsmcStorageValuation.Calculate, from host descriptors to host results
    shard = PathShard(rank, world, communicator_id(), total) if world > 1 else None
    valuation = LsmcStorageValuation(device=local_rank, shard=shard, reuse_handle=True)
    b_e2e, _ = build_params(cfg, total)
    params_e2e = b_e2e.Build()
    e2e_times = []
    for i in range(1 + max(1, args.e2e_steps)):  # the first call builds the tables and allocates; later calls re-use the handle
        barrier()
        t0 = time.perf_counter()
        res = valuation.Calculate(params_e2e)
        barrier()
        dt = time.perf_counter() - t0
        if i > 0:
            e2e_times.append(dt)
        if rank == 0:
            print(f"e2e call {i}: {dt:.3f} s (npv {res.Npv:.6f})", file=sys.stderr)
    e2e_cold_s = None
    valuation.Close()
    e2e_s = max_over_ranks(sorted(e2e_times)[len(e2e_times) // 2])
    d2h = 8 * (3 + (P + 1) * 6) + C.sizeof(_abi.lsmc_trigger_prices) * P + 8 * n_local
    h2d = sum(a.nbytes for a in flat.flat_storage.arrays.values()) + flat.forward.nbytes + flat.settle_df.nbytes + flat.powers.nbytes

    # ---- caller-supplied-simulation overload: panels in host memory, staged to the device inside the timed region
    e2e_panels = None
    if args.panel_sims > 0 and world == 1:
        e2e_panels = run_panels_overload(args, cfg, local_rank)

    # ---- N > 1: the weak-scaling figure (every rank owns the single-GPU path count) and sharded-vs-single parity
    weak = None
    parity = None
    if world > 1:
        if not args.no_weak:
            flat_w, h_w, n_w = open_handle(cfg, total * world)
            step_w, r_w, _ = timed_runs(h_w, max(1, min(2, args.steps)), 1)
            weak = {"value": lib.lsmc_num_evaluations(h_w.h) * world / step_w, "unit": UNIT, "ms_per_step": 1000.0 * step_w,
                    "paths_per_gpu": n_w, "total_paths": total * world, "steps": max(1, min(2, args.steps)), "warmup": 1}
            h_w.close()
        pc = dict(CONFIGS[3], sims=200_000, days=60, grid=100)
        _, h_s, _ = open_handle(pc, pc["sims"])
        r_s, arr_s = one_run(h_s, arrays=True)
        h_s.close()
        barrier()
        if rank == 0:
            _, h_1, _ = open_handle(pc, pc["sims"], sharded=False)
            r_1, arr_1 = one_run(h_1, arrays=True)
            h_1.close()
            dscale = np.abs(arr_1[0]).max()
            parity = {"case": "configs[2] storage, 200,000 paths x 60 daily steps x 100-point grid, same global path ids",
                      "npv_sharded": r_s.npv, "npv_single": r_1.npv, "npv_rel": abs(r_s.npv - r_1.npv) / abs(r_1.npv),
                      "std_error_rel": abs(r_s.std_error - r_1.std_error) / abs(r_1.std_error),
                      "delta_rel": float(np.abs(arr_s[0] - arr_1[0]).max() / dscale),
                      "profile_rel": float(np.abs(arr_s[1] - arr_1[1]).max() / np.abs(arr_1[1]).max())}
        barrier()

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return
    peak, peak_src = hbm_peak()
    # algorithmic bytes of the dominant kernel (fused backward step), per launch: every storage value written once and read
    # once (16 B per path*step*grid point) + one read of spot+factors for the basis of this and of the previous period
    mean_grid = evals_local / n_local / P
    F = cfg["factors"]
    bytes_per_launch = n_local * (16.0 * mean_grid + 2 * 8.0 * (1 + F))
    avg_kernel_s = max(acc["kernel_ms"], 1e-9) / 1000.0 / max(1, acc["bk_launches"])
    achieved = bytes_per_launch / avg_kernel_s / 1e9
    whole_bytes = n_local * P * (16.0 * mean_grid + 8.0 * (1 + F) * 3 + 32.0)  # SURVEY.md 8d B_alg, per GPU
    value = evals_total / step_s
    traffic, traffic_src = measured_traffic(args.config, n_local, cfg["grid"])
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": 1000.0 * step_s, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic", "config": workload_config(args.config, cfg, total, world),
           "evals_per_step": evals_total, "mean_grid_points": mean_grid, "npv": npv,
           "us_per_backward_step": 1000.0 * (phases["backward_ms"] - phases["gram_cholesky_ms"]) / P,
           "device_ms_per_step": acc["device_ms"] / args.steps, "phases_last_step": phases, "setup_seconds_untimed": setup_s,
           "roofline": {"bound": "hbm", "kernel": "backward_sweep_mma_kernel (dominant: one launch per backward step)", "achieved": achieved,
                        "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src, "traffic": traffic,
                        "traffic_source": traffic_src, "algorithmic_bytes_per_launch": bytes_per_launch,
                        "avg_launch_ms": 1000.0 * avg_kernel_s,
                        "timed_launches": int(acc["bk_launches"]), "launches_per_step": P,
                        "kernel_share_of_step": avg_kernel_s * P / step_s,
                        "whole_valuation": {"algorithmic_bytes_per_gpu": whole_bytes, "achieved": whole_bytes / step_s / 1e9,
                                            "frac": whole_bytes / step_s / 1e9 / peak}},
           "e2e": {"value": evals_total / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "seconds": e2e_s,
                   "what": "LsmcStorageValuation.Calculate (storage_b200/lsmc.py) from host descriptors to host results: device "
                           "simulation of both path sets + backward/forward + results copy (PvBySim included); the valuation object "
                           "keeps its handle between calls (reuse_handle), the first call, which builds tables and allocates, is "
                           "the warm-up; median of the --e2e-steps repetitions"},
           "gpu_launches": int(acc["launches"]), "clocks": clocks, "timed_seconds": total_s}
    if e2e_panels is not None:
        out["e2e_panels"] = e2e_panels
    if weak is not None:
        out["weak"] = weak
    if parity is not None:
        out["sharded_parity"] = parity
    if not args.no_cpu_baseline and world == 1:
        base, _, _ = cpu_baseline(CONFIGS[args.config], args.cpu_sims)
        out["cpu_baseline"] = base
    sys.stdout.flush()
    os.write(real_stdout, (json.dumps(out) + "\n").encode())
    if dist:
        dist.destroy_process_group()


def run_panels_overload(args, cfg, device):
    """LsmcStorageValuation.Calculate through UseSpotSimResults: both simulation sets are numpy panels in host memory."""
    from storage_b200 import FlatProblem, LsmcStorageValuation, SpotSimResultsFromPanels
    from tests import cases
    n = args.panel_sims
    b, model = build_params(cfg, n)
    probe = b.Build()
    flat = FlatProblem(probe, n, cfg["factors"])
    fwd = probe.ForwardCurve
    rng_rows = len(flat.sim_periods)
    t0 = time.perf_counter()
    reg_spot, reg_fac = cases.simulate_host(model, probe.CurrentPeriod, flat.sim_periods, fwd, n, 11)
    val_spot, val_fac = cases.simulate_host(model, probe.CurrentPeriod, flat.sim_periods, fwd, n, 23)
    gen_s = time.perf_counter() - t0
    b.UseSpotSimResults(SpotSimResultsFromPanels(flat.sim_periods, reg_spot, list(reg_fac)),
                        SpotSimResultsFromPanels(flat.sim_periods, val_spot, list(val_fac)))
    params = b.Build()
    valuation = LsmcStorageValuation(device=device, reuse_handle=True)
    times = []
    for i in range(3):
        t0 = time.perf_counter()
        res = valuation.Calculate(params)
        dt = time.perf_counter() - t0
        if i > 0:
            times.append(dt)
    valuation.Close()
    panel_bytes = 2 * (1 + cfg["factors"]) * rng_rows * n * 8
    from storage_b200 import _abi
    lib = _abi.load_library()
    evals = oracle_free_evals(lib, flat, n)
    dt = min(times)
    upload_ms = res.Timings.get("upload", 0.0)
    return {"value": evals / dt, "unit": UNIT, "seconds": dt, "paths": n, "h2d_bytes_per_step": int(panel_bytes), "npv": res.Npv,
            "h2d_gbs": panel_bytes / max(upload_ms, 1e-9) / 1e6 if upload_ms else None,
            "host_panel_generation_seconds_untimed": gen_s,
            "what": "UseSpotSimResults overload: 2 x (spot + %d factor) panels of %d rows x %d paths in pageable host memory, staged "
                    "through pinned buffers inside the timed call" % (cfg["factors"], rng_rows, n)}


def oracle_free_evals(lib, flat, n):
    """evaluations of one valuation from the library's own host tables (no GPU needed)."""
    P = flat.num_periods
    mn, mx = np.zeros(P), np.zeros(P)
    from storage_b200 import _abi
    from storage_b200.errors import check_status
    check_status(lib.lsmc_calculate_inventory_space(C.byref(flat.struct.storage), flat.struct.inventory, mn.ctypes.data_as(_abi.c_double_p),
                                                    mx.ctypes.data_as(_abi.c_double_p)), None)
    total = 1
    for p in range(1, P):
        total += lib.lsmc_grid_points(mn[p - 1], mx[p - 1], flat.struct.grid_spacing, None, 0)
    return total * n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=4, choices=sorted(CONFIGS), help="BASELINE.json shape, 1-based (SURVEY.md 8d)")
    ap.add_argument("--sims", type=int, default=0, help="override the configuration's TOTAL path count")
    ap.add_argument("--days", type=int, default=0, help="override the configuration's number of daily steps")
    ap.add_argument("--grid", type=int, default=0, help="override the configuration's inventory-grid points")
    ap.add_argument("--cpu-sims", type=int, default=4000, help="paths of the bounded CPU-baseline sample")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--panel-sims", type=int, default=0, help="paths of the caller-supplied-simulation e2e leg (0: skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the weak-scaling extra run")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        print(f"note: --warmup {args.warmup} < 3; the timing rules ask for at least 3 warm-up steps", file=sys.stderr)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
