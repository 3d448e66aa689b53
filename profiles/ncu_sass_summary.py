#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv` output: stall reasons, opcode mix, hottest SASS regions."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
for i, r in enumerate(data):  # ncu repeats the section once per view: keep the first
    if r and r[0] == "Kernel Name":
        data = data[:i]
        break
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
ops = collections.Counter()
ops_samples = collections.Counter()
n_inst = 0
for r in data:
    try:
        ex = int(r[col["Instructions Executed"]])
        smp = int(r[col["# Samples"]])
    except ValueError:
        continue
    n_inst += ex
    op = r[col["Source"]].split()[0].split(".")[0]
    if op.startswith("@"):
        op = r[col["Source"]].split()[1].split(".")[0]
    ops[op] += ex
    ops_samples[op] += smp
    for s in stall_cols:
        tot[s] += int(r[col[s]] or 0)
total_samples = sum(tot.values())
print("warp-level instructions executed:", n_inst, " SASS lines:", len(data))
print("stall reasons (share of samples):")
for s, v in tot.most_common(8):
    print(f"  {s:28s} {v:8d} {v / total_samples:.3f}")
print("opcode mix (share of executed instructions | share of samples):")
tot_s = sum(ops_samples.values())
for op, v in ops.most_common(22):
    print(f"  {op:12s} {v / n_inst:.3f} | {ops_samples[op] / tot_s:.3f}")
# hottest 12 instructions
hot = sorted(data, key=lambda r: -int(r[col["# Samples"]] or 0))[:14]
print("hottest instructions:")
for r in hot:
    print(f"  {r[col['# Samples']]:>6s}  {r[col['Source']].strip()[:90]}")
