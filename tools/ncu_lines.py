#!/usr/bin/env python
"""Per-source-line view of one kernel of an .ncu-rep: joins ncu's SASS page (per-address executed instructions, thread
instructions, stall samples) with nvdisasm's line info of the built library.
usage: ncu_lines.py <report.ncu-rep> <kernel regex> [launch index] [top N]"""
import collections
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_lines(kernel_mangled_substr):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "path_tracer_b200", "libptb.so")], cwd=tmp, capture_output=True)
    amap = {}
    for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
        out = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout
        cur, loc, inside = None, None, False
        for ln in out.splitlines():
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
            if m:
                inside = kernel_mangled_substr in m.group(1)
                cur = m.group(1)
                continue
            if not inside:
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
            if m:
                loc = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m and loc:
                amap[int(m.group(1), 16)] = loc
        if amap:
            break
    return amap


def main():
    rep, kre = sys.argv[1], sys.argv[2]
    launch = sys.argv[3] if len(sys.argv) > 3 else "0"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre, "--launch-skip", launch, "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    kname = rows[0][1]
    hdr = rows[1]
    ia, ii, it, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    # mangled-name substring: take the function name up to '<' or '('
    base = re.sub(r"^void ", "", kname).split("(")[0].split("<")[0].split("::")[-1]
    targs = re.search(r"<(.*?)>\(", kname)
    if targs:  # bool/int template arguments only (all this library uses): k<(bool)1, (bool)0> -> kILb1ELb0EE
        base = f"{len(base)}{base}I" + "".join(f"L{'b' if t == 'bool' else 'i'}{b}E" for t, b in re.findall(r"\((bool|int)\)(\d+)", targs.group(1))) + "E"
    else:
        base = f"{len(base)}{base}E"
    amap = sass_lines(base)
    base_addr = None
    agg = collections.defaultdict(lambda: [0, 0, 0])
    tot = [0, 0, 0]
    for r in rows[2:]:
        if len(r) <= it:
            if r and r[0] == "Kernel Name" and base_addr is not None:
                break  # next launch
            continue
        if r[ia] == "Address":
            continue
        a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
        if base_addr is None:
            base_addr = a
        loc = amap.get(a - base_addr, ("?", 0))
        v = [int(float(r[ii] or 0)), int(float(r[it] or 0)), int(float(r[isamp] or 0))]
        for k in range(3):
            agg[loc][k] += v[k]
            tot[k] += v[k]
    print(f"# {kname}\n# total warp-instr {tot[0]}, thread-instr {tot[1]} (avg active lanes {tot[1] / max(tot[0], 1):.2f}), samples {tot[2]}")
    print(f"{'file:line':28s} {'warp_inst%':>10s} {'lanes':>6s} {'samples%':>9s}")
    for loc, v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print(f"{loc[0] + ':' + str(loc[1]):28s} {100.0 * v[0] / tot[0]:10.2f} {v[1] / max(v[0], 1):6.2f} {100.0 * v[2] / max(tot[2], 1):9.2f}")
    # per file
    fa = collections.defaultdict(lambda: [0, 0, 0])
    for loc, v in agg.items():
        for k in range(3):
            fa[loc[0]][k] += v[k]
    print("# per file")
    for f, v in sorted(fa.items(), key=lambda x: -x[1][0]):
        print(f"{f:28s} {100.0 * v[0] / tot[0]:10.2f} {v[1] / max(v[0], 1):6.2f} {100.0 * v[2] / max(tot[2], 1):9.2f}")


if __name__ == "__main__":
    main()
