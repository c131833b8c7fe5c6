mkdir -p gpurun_out
rm -f gpurun_out/r3q_*.log
cat > /tmp/tl.py <<'PY'
import sys, os, json, time
sys.path.insert(0, os.getcwd())
import numpy as np
import path_tracer_b200 as ptb
from path_tracer_b200 import scenes
name, spp = sys.argv[1], int(sys.argv[2])
sd = {"c1": scenes.scene_c1, "c2": scenes.scene_c2, "c3": scenes.scene_c3, "c4": scenes.scene_c4, "c3s": lambda: scenes.scene_c3(grid=8)}[name]()
ctx = ptb.Context(0); ctx.upload(sd)
W, H = sd.width, sd.height
ctx.set_camera(sd.cam_origin, sd.cam_target, sd.fov_deg, sd.aspect)
ref = None
for tl in (0, 1, 0, 1):
    ctx.set_option("tile_lists", tl)
    ctx.frame_begin(W, H); ctx.render(0, 2, seed=1, max_bounces=sd.max_bounces)
    ms = []
    ctx.set_profiling(False)
    for rep in range(2):
        ctx.frame_begin(W, H); ctx.reset_counters()
        ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces); ctx.synchronize()
        ms.append(ctx.timing()["render"][0])
    a = ctx.read_accum(); pos, ids = ctx.read_aux(); c = ctx.counters()
    ctx.set_profiling(True); ctx.frame_begin(W, H); ctx.render(0, spp, seed=1, max_bounces=sd.max_bounces); ctx.synchronize(); t = ctx.timing(); ctx.set_profiling(False)
    key = (c["extend"], c["shadow"], c["probe"])
    if ref is None: ref = (a.copy(), pos.copy(), ids.copy(), key)
    na = int((a.view(np.uint32) != ref[0].view(np.uint32)).any(axis=2).sum()); npos = int((pos.view(np.uint32) != ref[1].view(np.uint32)).any(axis=2).sum()); nid = int((ids != ref[2]).sum())
    print(name, "tile_lists", tl, "ms", [round(x, 1) for x in ms], "pixels differing: accum", na, "position", npos, "id", nid, "counters equal", key == ref[3], key, "stages", {k: round(t[k][0], 1) for k in ("extend", "shade", "nee", "tail")}, flush=True)
PY
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "not full_size and not statistical" > gpurun_out/r3q_pytest.log 2>&1; tail -3 gpurun_out/r3q_pytest.log
for w in "c3 64" "c4 16"; do timeout 900 python /tmp/tl.py $w >> gpurun_out/r3q_tl.log 2>&1; done
cat gpurun_out/r3q_tl.log
