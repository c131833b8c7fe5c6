#!/usr/bin/env python
"""Basic-block view of one kernel of an .ncu-rep (SASS page): consecutive instructions with the same execution count form a block;
prints per block its size, executions, lanes per instruction, share of the warp instructions and of the stall samples, and its
instruction mix.   usage: ncu_blocks.py <report.ncu-rep> <kernel regex> [min share %]"""
import collections
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
min_share = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
out = subprocess.run(["ncu", "-i", rep, "--kernel-name", "regex:" + pat, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, L, kern = None, [], None
for r in rows:
    if not r:
        continue
    if r[0] == "Kernel Name":
        if kern is not None:
            break  # first matching launch only
        kern = r[1]
        continue
    if r[0] == "Address":
        hdr = r
        continue
    if hdr and len(r) >= len(hdr) - 2:
        d = dict(zip(hdr, r))
        L.append((d["Source"].strip(), float(d["Instructions Executed"]), float(d["Thread Instructions Executed"]), float(d["# Samples"])))
tot, tt, ts = sum(x[1] for x in L), sum(x[2] for x in L), sum(x[3] for x in L)
print(f"# {rep}: {kern}")
print(f"# {len(L)} instructions; {tot:.4g} warp instructions executed, {tt / tot:.2f} lanes per instruction, {int(ts)} stall samples")
blocks, cur = [], None
for i, x in enumerate(L):
    if cur and x[1] == cur[2]:
        cur[1] = i; cur[3] += x[2]; cur[4] += x[3]
    else:
        cur = [i, i, x[1], x[2], x[3]]
        blocks.append(cur)
print(f"{'instr range':>13s} {'n':>4s} {'executions':>11s} {'lanes':>6s} {'share%':>7s} {'samples%':>8s}  mix")
for b in blocks:
    n = b[1] - b[0] + 1
    share = 100 * n * b[2] / tot
    if share < min_share:
        continue
    ops = collections.Counter((x[0].split()[1] if x[0].startswith("@") else x[0].split()[0]).split(".")[0] for x in L[b[0]:b[1] + 1]).most_common(5)
    print(f"[{b[0]:5d}-{b[1]:5d}] {n:4d} {b[2]:11.4g} {b[3] / max(b[2] * n, 1):6.2f} {share:7.2f} {100 * b[4] / ts:8.2f}  " + " ".join(f"{k}:{v}" for k, v in ops))
