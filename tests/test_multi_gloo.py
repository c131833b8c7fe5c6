"""N > 1 path on CPU: world-size-2 gloo processes exercise the sample partition + the one reduce of the accumulation buffer
(path_tracer_b200/multi.py).  No GPU here, so each rank's producer is the oracle (the checker standing in for ptb_render); the
reduced image must equal the single-process render of the full sample range up to float summation order."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from path_tracer_b200.multi import sample_range  # noqa: E402


def test_sample_range_partitions_exactly():
    for begin, count, world in ((0, 64, 8), (5, 7, 4), (0, 3, 8), (100, 4096, 3), (0, 0, 2)):
        got = []
        for r in range(world):
            f, n = sample_range(begin, count, r, world)
            got += list(range(f, f + n))
        assert got == list(range(begin, begin + count))
        sizes = [sample_range(begin, count, r, world)[1] for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sample_range(0, 4, 2, 2)


def _worker(rank, world, port, W, H, begin, count, seed, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import po
    from path_tracer_b200 import scenes
    from path_tracer_b200.multi import render_partitioned
    sc = po.OracleScene(scenes.scene_c2())
    acc = torch.zeros(H * W * 4, dtype=torch.float32)

    def render_fn(first, n):
        r = sc.render(width=W, height=H, sample_begin=first, sample_count=n, seed=seed, threads=2)
        acc.add_(torch.from_numpy(np.ascontiguousarray(r["accum"]).reshape(-1)))

    render_partitioned(render_fn, acc, begin, count, rank, world)
    if rank == 0:
        np.save(out_path, acc.numpy().reshape(H, W, 4))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("count", [6, 5])
def test_two_rank_render_equals_single(tmp_path, count):
    import torch.multiprocessing as mp
    from oracle import po
    from path_tracer_b200 import scenes
    W, H, begin, seed, world = 64, 36, 3, 11, 2
    out = str(tmp_path / "acc.npy")
    port = 29500 + (os.getpid() % 2000) + count
    mp.spawn(_worker, args=(world, port, W, H, begin, count, seed, out), nprocs=world, join=True)
    got = np.load(out)
    ref = po.OracleScene(scenes.scene_c2()).render(width=W, height=H, sample_begin=begin, sample_count=count, seed=seed)["accum"]
    assert np.all(got[..., 3] == count)
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-6)  # identical sample set; only the float summation order differs
