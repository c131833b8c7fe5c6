mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2y_pytest.log
tail -4 gpurun_out/r2y_pytest.log
python bench.py > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err
python - <<'PY'
import json
r = json.loads(open("gpurun_out/r2y_bench.json").read().strip().splitlines()[-1])
print({k: r[k] for k in ("value", "ms_per_step", "e2e")}, r["roofline"].get("frac"), r["roofline"].get("stage_ms"))
for k, v in r.get("workloads", {}).items(): print(k, v.get("value"), v.get("ms_per_step"), v.get("roofline", {}).get("frac"))
print("pipelined", r.get("pipelined"))
PY
