"""Golden hashes of the build output (SURVEY Appendix A arrays) of the FULL-SIZE C4 scene (10 M-triangle terrain + 128 icospheres),
generated ONCE by the CPU oracle's literal restatement of blas_bvh.rs / tlas_bvh.rs / light_sampler.rs (oracle build ~40 s):
    python tests/golden/make_golden_c4.py
tests/test_build_parity.py::test_full_size_c4_build_golden_hash holds the product builder (scan-based, sub-tree parallel) to them."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def build_hashes(dumper, n_blas):
    """sha256 over the raw bytes of every Appendix-A array, per BLAS, per TLAS and for the light sampler."""
    def h(*arrays):
        m = hashlib.sha256()
        for a in arrays:
            m.update(np.ascontiguousarray(a).tobytes())
        return m.hexdigest()

    out = {}
    for b in range(n_blas):
        d = dumper.blas_dump(b)
        out[f"blas{b}"] = dict(nodes=int(d["meta"].shape[0]), root=int(d["root"]), meta=h(d["meta"]), boxes=h(d["boxes"].view(np.uint32)), leaf_prims=h(d["leaf_prims"]),
                               triangles=h(d["triangles"].view(np.uint32)))
    for which in (0, 1):
        d = dumper.tlas_dump(which)
        out[f"tlas{which}"] = dict(nodes=int(d["meta"].shape[0]), root=int(d["root"]), meta=h(d["meta"]), boxes=h(d["boxes"].view(np.uint32)), matrices=h(d["matrices"].view(np.uint32)),
                                   blas_ids=h(d["blas_ids"]))
    ls = dumper.light_sampler()
    out["lights"] = dict(n=int(ls["items"].shape[0]), items=h(ls["items"]), pdf=h(ls["pdf"].view(np.uint32)), cdf=h(ls["cdf"].view(np.uint32)),
                         max_bits=int(np.float32(ls["max"]).view(np.uint32)))
    return out


if __name__ == "__main__":
    from oracle import po
    from path_tracer_b200 import scenes
    sd = scenes.scene_c4()
    sc = po.OracleScene(sd)
    hashes = dict(scene=sd.name, triangles=int(sd.verts.shape[0]), generator="oracle/po_accel.hpp via tests/golden/make_golden_c4.py", arrays=build_hashes(sc, len(sd.blas)))
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "build_c4_full_hashes.json"), "w") as f:
        json.dump(hashes, f, indent=1)
    print("written:", hashes["triangles"], "triangles,", hashes["arrays"]["blas1"]["nodes"] if "blas1" in hashes["arrays"] else "", "nodes")
