"""Generates the committed golden vectors from the CPU oracle (run in the authoring container: python tests/golden/make_golden.py).
The reference ships no fixtures of its own, so these pin the oracle's outputs at the time of writing."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import po  # noqa: E402
from path_tracer_b200 import scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
idx = (np.arange(4096, dtype=np.uint32) * 7) % 5
seed = (np.arange(4096, dtype=np.uint32) * np.uint32(2654435761)).astype(np.uint32)
np.savez_compressed(os.path.join(HERE, "integer_tier.npz"), index=idx, seed=seed, jitter_bits=po.ss_sobol(idx, seed).view(np.uint32), sobol_table=po.sobol_table(),
                    rng_77_12345_9=po.rng_stream(77, 12345, 9, 64))
sd = scenes.scene_c2()
sc = po.OracleScene(sd)
o, d, tm = scenes.incoherent_rays(sd, 8192, seed=4242)
_, _, tf = scenes.incoherent_rays(sd, 8192, seed=4242, finite_tmax=True)
r = sc.trace_closest(o, d, tm)
np.savez_compressed(os.path.join(HERE, "trace_c2.npz"), o=o, d=d, tmax=tm, tmax_finite=tf, hit=r["hit"], prim=r["prim"], inst=r["inst"], t_bits=r["t"].view(np.uint32),
                    any_hit=sc.trace_any(o, d, tf)["hit"])
d0 = sc.blas_dump(5)
np.savez_compressed(os.path.join(HERE, "build_c2.npz"), meta=d0["meta"], boxes_bits=d0["boxes"].view(np.uint32), leaf_prims=d0["leaf_prims"], tlas_meta=sc.tlas_dump(0)["meta"],
                    tlas_boxes_bits=sc.tlas_dump(0)["boxes"].view(np.uint32))
im = sc.render(width=64, height=36, sample_count=4, seed=5)
np.savez_compressed(os.path.join(HERE, "render_c2_64x36_4spp_seed5.npz"), accum=im["accum"], id=im["id"])
# State::update sequence (5 frames: static, moved, moved, static, moved) on seeded inputs: accumulation texture bits after every frame
sys.path.insert(0, os.path.join(ROOT, "tests"))
import temporal_inputs  # noqa: E402


class _OracleState:
    def __init__(self, W, H):
        self.acc = np.zeros((H, W, 4), np.float32)
        self.vel = np.zeros((H, W, 2), np.float32)

    def update(self, data, pos, ids, inv, last):
        self.acc, vel, ran = po.state_update(data, pos, ids, inv, last, self.acc)
        if ran:
            self.vel = vel


stt = _OracleState(48, 27)
seq = temporal_inputs.run_sequence(stt.update, lambda: stt.acc.copy(), po, 48, 27, 5, seed=11)
np.savez_compressed(os.path.join(HERE, "temporal_48x27_seed11.npz"), acc_bits=np.stack(seq).view(np.uint32), vel_bits=stt.vel.view(np.uint32))
print("golden vectors written")
