#!/usr/bin/env python
"""bench.py — LSMC path*step*inventory-point evaluations per second on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA library through the C ABI)
  python bench.py --impl reference --steps K --warmup W    the reference algorithm's CPU port (oracle) on the host cores

A "step" is one full pass of the hot path — backward induction + forward simulation (lsmc_run) — over one batch of
simulated paths already resident in HBM.  Workload at N=1: BASELINE.json configs[3] shape (5-year daily, 3-factor seasonal
model, 1M paths, 200-point inventory grid, ratchets with outages, loss, consumed commodity, inventory cost, empty at
end) on ONE GPU; with N>1 every rank owns the same number of paths (weak scaling) and the ranks exchange the per-step
regression sums over NCCL.  `e2e` is the same metric through the public Calculate call from host descriptors: table
build, device simulation of both path sets, the run, and the device->host copy of the results, all timed.
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


def build_params(num_sims, num_days, grid_points, seed=11):
    """LsmcValuationParameters for the bench workload, built through the host API mirror (built-in simulation overload)."""
    from storage_b200 import (BasisFunctionsBuilder, Day, FixedSpacingStateSpaceGridCalc, LsmcValuationParameters,
                              SimulationDataReturned, StorageHelper, settle_on_20th_of_following_month)
    from tests import cases
    start = Day(2020, 4, 1)
    end = start + num_days
    val_date = start - 1
    storage = cases.ratchet_storage(start, end, outages=True, kind="linear")
    model = cases.three_factor_model(start, end)
    b = LsmcValuationParameters.Builder()
    b.CurrentPeriod = val_date
    b.Inventory = 5_250.0  # 10 % of the maximum inventory
    b.ForwardCurve = cases.forward_curve(val_date, end)
    b.Storage = storage
    b.SettleDateRule = settle_on_20th_of_following_month
    b.DiscountFactors = StorageHelper.CreateAct65ContCompDiscounter(0.055)
    b.GridCalc = FixedSpacingStateSpaceGridCalc.CreateForFixedNumberOfPointsOnGlobalInventoryRange(storage, grid_points)
    b.BasisFunctions = BasisFunctionsBuilder.Parse(cases.THREE_FACTOR_BASIS_TEXT, 3)
    b.SimulationDataReturned = SimulationDataReturned.Nothing
    b.SimulateWithMultiFactorModelAndMersenneTwister(model, num_sims, seed, seed * 2)
    return b.Build(), model


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


def cpu_baseline(args, flat_shape):
    """The oracle (port of the reference algorithm, QR regression in double, OpenMP over grid points / paths) on a bounded
    sample of the same workload shape."""
    from oracle import oracle
    from storage_b200 import FlatProblem
    from tests import cases
    num_days, grid_points = flat_shape
    n = args.cpu_sims
    params, model = build_params(n, num_days, grid_points)
    flat = FlatProblem(params, n, 3)
    sim_periods = flat.sim_periods
    fwd = params.ForwardCurve
    reg_spot, reg_fac = cases.simulate_host(model, params.CurrentPeriod, sim_periods, fwd, n, 11)
    val_spot, val_fac = cases.simulate_host(model, params.CurrentPeriod, sim_periods, fwd, n, 23)
    evals = sum(len(oracle_grid) for oracle_grid in grids_of(flat)) * n
    threads = oracle.num_threads()
    t0 = time.perf_counter()
    res = oracle.lsmc_calculate(flat, reg_spot, list(reg_fac), val_spot, list(val_fac), mode=0, threads=threads, panels=False)
    dt = time.perf_counter() - t0
    return {"value": evals / dt, "unit": "path*step*inventory-point evals/s", "cores": threads, "kind": "port",
            "sample": f"{n} paths x {flat.num_periods} daily steps x {grid_points}-point grid, same storage/model/basis as the "
                      f"GPU workload, QR regression in double, {dt:.1f} s",
            "seconds": dt, "npv": res.npv}, evals, dt


def grids_of(flat):
    """Per-period inventory grids (decision periods only), from the library's host code (no GPU needed)."""
    from storage_b200 import _abi
    lib = _abi.load_library()
    P = flat.num_periods
    mn, mx = np.zeros(P), np.zeros(P)
    st = flat.struct.storage
    from storage_b200.errors import check_status
    check_status(lib.lsmc_calculate_inventory_space(C.byref(st), flat.struct.inventory, mn.ctypes.data_as(_abi.c_double_p),
                                                    mx.ctypes.data_as(_abi.c_double_p)), None)
    grids = [[flat.struct.inventory]]
    for p in range(1, P):
        n = lib.lsmc_grid_points(mn[p - 1], mx[p - 1], flat.struct.grid_spacing, None, 0)
        grids.append(range(n))
    return grids


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU reference runs on rank 0 only
    shape = (args.days, args.grid)
    times, value = [], None
    for i in range(args.warmup + args.steps):
        base, evals, dt = cpu_baseline(args, shape)
        if i >= args.warmup:
            times.append(dt)
    ms = 1000.0 * sum(times) / len(times)
    value = evals / (ms / 1000.0)
    base["value"] = value
    out = {"impl": "reference", "metric": "LSMC path*step*inventory-point evals/sec", "value": value,
           "unit": "path*step*inventory-point evals/s", "n_gpus": 0, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": workload_config(args, args.cpu_sims, 1, cpu=True),
           "cpu_baseline": base,
           "e2e": {"value": value, "unit": "path*step*inventory-point evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def workload_config(args, sims_per_gpu, world, cpu=False):
    return {"workload": "BASELINE.json configs[3] shape: %d daily steps, 3-factor seasonal model, basis 1+x0+x0^2+x1+x2+x0*x1+s, "
                        "%d-point inventory grid, time/inventory-varying ratchets with outages, loss, consumed cmdty, inventory cost, "
                        "empty at end" % (args.days, args.grid),
            "paths_per_gpu": sims_per_gpu, "total_paths": sims_per_gpu * world, "steps_per_valuation": args.days,
            "grid_points": args.grid, "factors": 3, "basis_functions": 7,
            "parallelism": ("cpu-openmp" if cpu else f"path-sharded x{world}, per-step NCCL all-gather of regression sums"),
            "cache": "inputs (simulated paths, >100 GB) and the per-step value slabs (3.2 GB) are far larger than the 126 MB L2"}


def run_gpu(args):
    import torch
    from storage_b200 import FlatProblem, Handle, _abi
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: storage_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n = args.sims
    params, model = build_params(n, args.days, args.grid)
    lib = _abi.load_library()

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    comm_id = {}

    def join_ranks(handle):
        """Rank 0 draws a communicator id once, torch.distributed carries it to the other ranks; every handle then joins
        with the same id (the library creates the NCCL communicator on first use and shares it afterwards)."""
        if world == 1:
            return
        if "raw" not in comm_id:
            ident = torch.zeros(_abi.LSMC_COMM_ID_BYTES, dtype=torch.uint8)
            if rank == 0:
                buf = (C.c_ubyte * _abi.LSMC_COMM_ID_BYTES)()
                handle.check(lib.lsmc_comm_unique_id(buf))
                ident = torch.tensor(list(buf), dtype=torch.uint8)
            ident = ident.cuda()
            dist.broadcast(ident, 0)
            comm_id["raw"] = bytes(ident.cpu().tolist())
        handle.check(lib.lsmc_comm_init(handle.h, comm_id["raw"], rank, world, n * world))

    def open_handle():
        """Everything a Calculate call does before the run: descriptor flattening, table build, device buffers, simulation."""
        flat_ = FlatProblem(params, n, 3)
        handle = Handle(flat_, local_rank)
        join_ranks(handle)
        mp, keep = flat_.model_arrays(model, params.CurrentPeriod)
        handle.check(lib.lsmc_simulate(handle.h, 0, C.byref(mp), 11, 1, rank * n))
        handle.check(lib.lsmc_simulate(handle.h, 1, C.byref(mp), 22, 1, rank * n))
        del keep
        return flat_, handle

    def one_run(handle, full_results=False):
        handle.check(lib.lsmc_run(handle.h, None, None, None))
        r = _abi.lsmc_results()
        keep = None
        if full_results:
            P_ = handle.flat.num_periods
            keep = (np.zeros(P_ + 1), np.zeros((P_ + 1, 5)), (_abi.lsmc_trigger_prices * P_)())
            r.deltas = keep[0].ctypes.data_as(_abi.c_double_p)
            r.profile = keep[1].ctypes.data_as(_abi.c_double_p)
            r.trigger_prices = keep[2]
        handle.check(lib.lsmc_get_results(handle.h, C.byref(r)))
        return r

    # ---- resident-input arm: one handle, simulations generated on the device before the timed region
    t_setup = time.perf_counter()
    flat, h = open_handle()
    setup_s = time.perf_counter() - t_setup
    evals_local = lib.lsmc_num_evaluations(h.h)
    evals_total = evals_local * world

    for _ in range(args.warmup):
        one_run(h)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    times, kernel_ms, launches, bk_launches, device_ms = [], 0.0, 0, 0, 0.0
    t_all = time.perf_counter()
    for _ in range(args.steps):
        barrier()
        t0 = time.perf_counter()
        r = one_run(h)
        barrier()
        times.append(time.perf_counter() - t0)
        phases = {"gram_cholesky_ms": r.ms_pseudo_inverse, "backward_ms": r.ms_backward, "forward_ms": r.ms_forward,
                  "run_total_ms": r.ms_total}
        kernel_ms += r.ms_backward_kernel
        bk_launches += r.backward_kernel_launches
        launches += r.kernel_launches
        device_ms += r.ms_device
    total_s = time.perf_counter() - t_all
    clocks = sampler.stop() if rank == 0 else None
    npv = r.npv
    step_s = sum(times) / len(times)
    if dist:
        t = torch.tensor([step_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        step_s = float(t.item())
    h.close()

    # ---- end-to-end arm: the whole Calculate sequence from host descriptors, per step: table build, device buffers, the
    # (rank-joining,) device simulation of both path sets, the run, and the device->host copy of all scalar results
    e2e_times = []
    for i in range(1 + max(1, args.e2e_steps)):
        barrier()
        t0 = time.perf_counter()
        f2, h2 = open_handle()
        t1 = time.perf_counter()
        one_run(h2, full_results=True)
        t2 = time.perf_counter()
        h2.close()
        barrier()
        if i > 0:
            e2e_times.append(time.perf_counter() - t0)
            if rank == 0:
                print(f"e2e step {i}: open {t1 - t0:.3f} s, run+results {t2 - t1:.3f} s, close {time.perf_counter() - t2:.3f} s", file=sys.stderr)
    d2h = 8 * (3 + (flat.num_periods + 1) * 6) + C.sizeof(_abi.lsmc_trigger_prices) * flat.num_periods
    h2d = sum(a.nbytes for a in flat.flat_storage.arrays.values()) + flat.forward.nbytes + flat.settle_df.nbytes + flat.powers.nbytes
    e2e_s = sorted(e2e_times)[len(e2e_times) // 2]  # median: one cudaMalloc/cudaFree hiccup of the 120 GB working set is not the metric
    if dist:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return
    peak, peak_src = hbm_peak()
    # algorithmic bytes of the dominant kernel (fused backward step), per launch: every storage value written once and
    # read once (16 B per path*step*grid point) + one read of spot+factors for the basis of this and of the previous period
    P = flat.num_periods
    mean_grid = evals_local / n / P
    bytes_per_launch = n * (16.0 * mean_grid + 2 * 8.0 * (1 + 3))
    avg_kernel_s = max(kernel_ms, 1e-9) / 1000.0 / max(1, bk_launches)
    achieved = bytes_per_launch / avg_kernel_s / 1e9
    value = evals_total / step_s
    out = {"metric": "LSMC path*step*inventory-point evals/sec", "value": value, "unit": "path*step*inventory-point evals/s",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * step_s, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": workload_config(args, n, world),
           "evals_per_step": evals_total, "mean_grid_points": mean_grid, "npv": npv,
           "device_ms_per_step": device_ms / args.steps, "phases_last_step": phases, "setup_seconds_untimed": setup_s,
           "roofline": {"bound": "hbm", "kernel": "backward_sweep_ring_kernel<7,fused>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "peak_source": peak_src,
                        # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch at a full-width step (1M paths, 200 grid
                        # points; profiles/r01_backward_sweep_ring.txt), scaled to this run's paths and mean grid width
                        "traffic": 3.376e9 * (n / 1.0e6) * (16.0 * mean_grid + 64.0) / (16.0 * 200 + 64.0),
                        "traffic_source": "ncu --set full, one launch, profiles/r01_backward_sweep_ring.txt (1.034 x algorithmic)",
                        "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_ms": 1000.0 * avg_kernel_s,
                        "kernel_share_of_step": (kernel_ms / 1000.0 / args.steps) / step_s},
           "e2e": {"value": evals_total / e2e_s, "unit": "path*step*inventory-point evals/s", "h2d_bytes_per_step": int(h2d),
                   "d2h_bytes_per_step": int(d2h), "seconds": e2e_s,
                   "what": "LsmcStorageValuation.Calculate from host descriptors: table build + device simulation of both path sets "
                           "+ backward/forward + results copy + handle teardown; median of the --e2e-steps repetitions"},
           "gpu_launches": int(launches), "clocks": clocks, "timed_seconds": total_s}
    if not args.no_cpu_baseline and world == 1:
        base, _, _ = cpu_baseline(args, (args.days, args.grid))
        out["cpu_baseline"] = base
    print(json.dumps(out))
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sims", type=int, default=1_000_000, help="paths per GPU")
    ap.add_argument("--days", type=int, default=1825)
    ap.add_argument("--grid", type=int, default=200)
    ap.add_argument("--cpu-sims", type=int, default=4000, help="paths of the bounded CPU-baseline sample")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        print(f"note: --warmup {args.warmup} < 3; the timing rules ask for at least 3 warm-up steps", file=sys.stderr)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
