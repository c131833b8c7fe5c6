mkdir -p gpurun_out
for w in c2 c3 c4; do python bench.py --workload $w > gpurun_out/g2_bench_$w.json 2> gpurun_out/g2_bench_$w.err; done
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python - <<'PY'
import json
for w in ('c2','c3','c4'):
    j=json.load(open(f'gpurun_out/g2_bench_{w}.json')); print(w, round(j['value'],1), 'e2e', round(j['e2e']['value'],1), 'ms', round(j['ms_per_step'],2), 'frac', round(j['roofline']['frac'],3))
PY
