"""profiles/ncu_counters.json (machine-readable hardware counters of the dominant kernel, read by bench.py into roofline.ncu) and
profiles/traffic.json from kernel summaries written by tools/ncu_summary.py report:
python tools/update_ncu_counters.py r2m "note"      (reads profiles/<tag>_c{2,3,4}_kernels_ncu.txt)"""
import json
import re
import subprocess
import sys

tag, note = sys.argv[1], sys.argv[2]
out = {}
for wl in ("c2", "c3", "c4"):
    path = f"profiles/{tag}_{wl}_kernels_ncu.txt"
    txt = open(path).read()

    def row(key):
        m = re.search(r"^" + re.escape(key) + r" \[([^\]]*)\]: (.*)$", txt, re.M)
        if not m:
            return None, None
        return m.group(1), [float(v.replace(",", "")) for v in m.group(2).split()]

    kern = re.search(r"^kernel: (.*)$", txt, re.M).group(1)
    names = [k.strip().strip("'") for k in kern.strip("[]").split("', '")]
    # the dominant traversal launch: k_extend_mr (continuing slots, two rays per lane) when the scene uses it, else k_extend
    col = next((i for i, k in enumerate(names) if "k_extend_mr" in k), next(i for i, k in enumerate(names) if "k_extend" in k))
    rec = dict(source=path, kernels=kern, captured_launch=f"{names[col]} = column {col}: an early, full-size wavefront iteration (wavefront_groups = 1)", note=note)
    for name, key in (("issue_active", "smsp__issue_active.avg.pct_of_peak_sustained_active"), ("lanes_per_inst", "smsp__thread_inst_executed_per_inst_executed.ratio"),
                      ("dram_frac", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), ("l2_frac", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
                      ("alu_pipe", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"), ("fma_pipe", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                      ("warps_active", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("l1_hit", "l1tex__t_sector_hit_rate.pct"), ("l2_hit", "lts__t_sector_hit_rate.pct")):
        unit, v = row(key)
        if v:
            rec[name] = v[col] / 100.0 if unit == "%" else v[col]
    unit, v = row("gpu__time_duration.sum")
    rec["launch_ms"] = v[col] * {"ms": 1.0, "us": 1e-3, "s": 1e3}.get(unit, 1.0)
    rec["issue_x_lanes"] = rec["issue_active"] * rec["lanes_per_inst"] / 32.0  # fraction of the SIMT issue x lane capacity in use
    out[wl] = rec
    subprocess.check_call([sys.executable, "tools/update_traffic.py", wl, path,
                           "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of ONE captured launch of each kernel (an early, full-size wavefront iteration, wavefront_groups = 1, "
                           "16 slots per pixel: larger than the average launch of a render); includes the path-state records the kernels read and write"])
json.dump(out, open("profiles/ncu_counters.json", "w"), indent=1)
print(json.dumps(out, indent=1))
