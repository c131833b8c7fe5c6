mkdir -p gpurun_out
rm -f gpurun_out/r3b_*.log
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "multiray or same_streams or groups or path_kernel or ordered or schedule or furnace" > gpurun_out/r3b_pytest.log 2>&1; tail -3 gpurun_out/r3b_pytest.log
cat > /tmp/ov.py <<'PY'
import sys, os, json, time
sys.path.insert(0, os.getcwd())
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes
name, spp = sys.argv[1], int(sys.argv[2])
sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4}[name]()
ctx = ptb.Context(0); ctx.upload(sd)
W, H = sd.width, sd.height
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
ref = None
for ov in (0, 4, 0, 1, 2, 4, 8, 16, -1):
    ctx.set_option("nee_overlap", ov)
    ctx.frame_begin(W, H); ctx.render(0, 2, seed=1, max_bounces=sd.max_bounces)
    ms = []
    for rep in range(3):
        ctx.frame_begin(W, H); ctx.reset_counters()
        ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces); ctx.synchronize()
        ms.append(ctx.timing()["render"][0]); ctx.set_profiling(False)
    a = ctx.read_accum(); c = ctx.counters()
    key = (a.view(np.uint32).sum(dtype=np.uint64), c["extend"], c["shadow"], c["probe"])
    if ref is None: ref = (a.copy(), key)
    same = bool(np.array_equal(a.view(np.uint32), ref[0].view(np.uint32))) and key == ref[1]
    print(name, "nee_overlap", ov, "ms", [round(x, 1) for x in ms], "identical", same, flush=True)
PY
for w in "c3 64" "c4 16" "c2 256" "c1 64"; do timeout 900 python /tmp/ov.py $w >> gpurun_out/r3b_overlap.log 2>&1; done
cat gpurun_out/r3b_overlap.log
