#!/usr/bin/env python
"""Small valuations through every backward-kernel variant, the streamed forward pass and the device simulator, for
compute-sanitizer (memcheck / racecheck / synccheck), e.g.
    compute-sanitizer --tool racecheck python profiles/sanitizer_case.py
Results are compared with the default variant's, so a tool that perturbs timing also shows up as a numerical difference."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from storage_b200 import LsmcStorageValuation, SimulationDataReturned  # noqa: E402
from tests import cases  # noqa: E402

VARIANTS = {"mma": {}, "ring_dfma": {"LSMC_SWEEP_MMA": "0"}, "direct_sweep": {"LSMC_SWEEP_RING": "0"}, "general": {"LSMC_SWEEP": "0"},
            "mma_pieces": {"LSMC_MMA_TAIL_SPLIT": "2,3"}}
KEYS = sorted({k for v in VARIANTS.values() for k in v})
base = None
for name, env in VARIANTS.items():
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update(env)
    for wide in (False, True):
        if wide:
            from storage_b200 import BasisFunctionsBuilder, Sim
            basis = (BasisFunctionsBuilder.Ones + Sim.X0 + Sim.X0.Pow(2) + Sim.X1 + Sim.X2 + Sim.X0 * Sim.X1 + Sim.S + Sim.S.Pow(2)
                     + Sim.X1.Pow(2) + Sim.X2.Pow(2))
            case = cases.three_factor_ratchets(num_sims=700, num_days=12, seed=17, grid_points=40, basis=basis)
        else:
            case = cases.three_factor_ratchets(num_sims=700, num_days=12, seed=13, grid_points=40)
        for order in (False, True):
            res = LsmcStorageValuation().Calculate(case.params, reference_operation_order=order)
            key = (wide,)
            if name == "mma" and not order:
                base = base or {}
                base[key] = res.Npv
            rel = abs(res.Npv - base[key]) / abs(base[key])
            print(f"{name:13s} K={'10' if wide else ' 7'} order={'reference' if order else 'fused    '} npv {res.Npv:.9f} rel diff {rel:.1e}")
            assert rel < 1e-9
for k in KEYS:
    os.environ.pop(k, None)
# streamed panels + device simulator
os.environ["LSMC_STREAM_BLOCK"] = "5"
case = cases.simple_one_factor(num_sims=700, num_days=12, seed=3, flags=SimulationDataReturned.All)
a = LsmcStorageValuation().Calculate(case.params, keep_decisions=True, stream_panels=True)
b = LsmcStorageValuation().Calculate(case.params, keep_decisions=True, stream_panels=False)
assert a.Npv == b.Npv and (a.InventoryBySim == b.InventoryBySim).all()
bm = case.builder.Clone()
bm.SimulateWithMultiFactorModelAndMersenneTwister(case.model, 700, 7, 8)
print("device simulator npv", LsmcStorageValuation().Calculate(bm.Build()).Npv)
# determinism stress of the mbarrier hand-offs: a race in the row ring or in the regression-sum hand-off would change some value
# in some run; every run must repeat the first to the bit (NPV, deltas, every path's PV)
import numpy as np  # noqa: E402
os.environ.pop("LSMC_STREAM_BLOCK", None)
case = cases.three_factor_ratchets(num_sims=60_000, num_days=30, seed=5, grid_points=200)
first = None
for i in range(12):
    r = LsmcStorageValuation().Calculate(case.params)
    sig = (r.Npv, tuple(r.Deltas.values()), hash(r.PvBySim.tobytes()))
    first = first or sig
    assert sig == first, f"run {i} differs from run 0"
print("12 runs of 60,000 paths x 30 days x 200 grid points: bit-identical")
print("SANITIZER_CASE_OK")
