"""profiles/traffic.json entry of one workload from a kernels summary written by tools/ncu_summary.py report:
python tools/update_traffic.py c2 profiles/r1m_c2_kernels_ncu.txt "note" """
import json, re, sys
wl, path, note = sys.argv[1], sys.argv[2], sys.argv[3]
txt = open(path).read()
def row(key):
    m = re.search(r"^" + re.escape(key) + r" \[(\w+)\]: (.*)$", txt, re.M)
    unit, vals = m.group(1), [float(v) for v in m.group(2).split()]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3}[unit]
    return [v * scale for v in vals]
kern = re.search(r"^kernel: (.*)$", txt, re.M).group(1)
rd, wr, dur = row("dram__bytes_read.sum"), row("dram__bytes_write.sum"), row("gpu__time_duration.sum")
tot = [a + b for a, b in zip(rd, wr)]
p = "profiles/traffic.json"
j = json.load(open(p))
names = [k.strip().strip("'") for k in kern.strip("[]").split("', '")]
def col(prefix, skip=()):  # first captured launch of a kernel by name (an iteration is three or, with the multiray kernels, four launches)
    for i, k in enumerate(names):
        if prefix in k and not any(x in k for x in skip):
            return i
    return None
ie = col("k_extend")
ish, ine = col("k_shade"), col("k_nee")
j[wl] = {"k_extend_dram_bytes_per_launch": tot[ie], "k_shade_dram_bytes_per_launch": tot[ish], "k_nee_dram_bytes_per_launch": tot[ine],
         "captured_launch_duration": " / ".join(f"{d:.6g} ms" for d in dur), "kernels": kern, "source": path, "note": note}
ie2 = [i for i, k in enumerate(names) if "k_extend" in k and i != ie]
if ie2:
    j[wl]["k_extend_part2_dram_bytes_per_launch"] = tot[ie2[0]]
    j[wl]["note_parts"] = "k_extend is launched in two parts with the multiray kernels: continuing slots (k_extend_mr, the column reported as k_extend) and slots to regenerate (k_extend<7>, part 2)"
json.dump(j, open(p, "w"), indent=1)
print(j[wl])
