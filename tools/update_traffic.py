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
j[wl] = {"k_extend_dram_bytes_per_launch": tot[0], "k_shade_dram_bytes_per_launch": tot[1], "k_nee_dram_bytes_per_launch": tot[2],
         "captured_launch_duration": " / ".join(f"{d:.6g} ms" for d in dur), "kernels": kern, "source": path, "note": note}
json.dump(j, open(p, "w"), indent=1)
print(j[wl])
