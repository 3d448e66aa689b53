#!/usr/bin/env python
"""Regenerates tests/golden/oracle_small_cases.json: results of the CPU oracle (both regression modes) on three small
seeded cases.  The reference's own literal goldens (test_multi_factor.py) need the .NET Mersenne-Twister stream and cannot
be reproduced in this image, so these fixtures pin the ORACLE against regressions, not the oracle against the reference
(see DESIGN.md section 2).  Run from the repo root:  python tests/golden/make_golden.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402
from tests import cases  # noqa: E402

CASES = {
    "one_factor_cubic_60d_500": lambda: cases.simple_one_factor(num_sims=500, num_days=60, seed=11),
    "three_factor_ratchets_45d_400": lambda: cases.three_factor_ratchets(num_sims=400, num_days=45, seed=4),
    "active_period_terminal_value_20d_300": lambda: cases.simple_one_factor(num_sims=300, num_days=20, seed=6, val_offset=4,
                                                                            terminal=(-999.0, 0.0, 0.0, 1.0)),
}


def main():
    out = {}
    for name, make in CASES.items():
        case = make()
        entry = {}
        for mode, label in ((0, "qr"), (1, "exact")):
            r = oracle.lsmc_calculate(case.flat, case.reg_spot, case.reg_factors, case.val_spot, case.val_factors, mode=mode)
            entry[label] = {"npv": r.npv, "std_error": r.std_error, "backward_npv": r.backward_npv,
                            "deltas": list(r.deltas), "profile_inventory": list(r.profile[:, 0]),
                            "decision_checksum": int(r.decision_index.sum()),
                            "pv_first": list(r.pv_by_sim[:5])}
        out[name] = entry
    path = os.path.join(ROOT, "tests", "golden", "oracle_small_cases.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path)


if __name__ == "__main__":
    main()
