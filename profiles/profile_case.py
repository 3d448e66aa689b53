#!/usr/bin/env python
"""One valuation of a bench-shaped workload (built-in device simulation), for ncu captures:
   python profiles/profile_case.py --sims 1000000 --days 120 --grid 200 [--runs 1]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--sims", type=int, default=1_000_000)
ap.add_argument("--days", type=int, default=120)
ap.add_argument("--grid", type=int, default=200)
ap.add_argument("--runs", type=int, default=1)
a = ap.parse_args()
from storage_b200 import LsmcStorageValuation  # noqa: E402
cfg = dict(bench.CONFIGS[4], sims=a.sims, days=a.days, grid=a.grid)
builder, _ = bench.build_params(cfg, a.sims)
params = builder.Build()
for i in range(a.runs):
    t0 = time.perf_counter()
    res = LsmcStorageValuation().Calculate(params)
    print(f"run {i}: npv {res.Npv:.6f}  {time.perf_counter() - t0:.3f} s  timings {res.Timings}")
