"""Sample-partitioned rendering across the GPUs of one box (one process per GPU, torch.distributed).

The reference is a single process on CPU threads; its frame loop (main.rs:181-207) is a pure map over (pixel, sample) and the
accumulation (accumulate.wgsl:23) a commutative float sum, so the path shards by *global sample index*: every rank holds the
full scene, renders a disjoint sample range for all pixels and the W x H x 4 accumulation buffers are summed onto rank 0 with
ONE collective per frame batch (NCCL reduce over NVLink on the GPU box; gloo in the CPU tests).  RNG streams are keyed by
(seed, pixel, global sample index), so the union of the samples is the same for every world size.
"""
from __future__ import annotations


def sample_range(begin: int, count: int, rank: int, world: int):
    """Contiguous block of the global sample range [begin, begin+count) owned by `rank`: the first count % world ranks get one
    sample more.  Returns (first global sample index, number of samples); count may be 0 for ranks beyond the sample count."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(count, world)
    mine = base + (1 if rank < extra else 0)
    first = begin + rank * base + min(rank, extra)
    return first, mine


def reduce_accum(acc, dst: int = 0, group=None):
    """Sum the accumulation buffers (torch tensor, W*H*4 float32: sum rgb, sample count) of all ranks onto `dst`, in place."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return acc


def render_partitioned(render_fn, acc, begin: int, count: int, rank: int, world: int, dst: int = 0, group=None):
    """One frame batch on `world` ranks: `render_fn(first, n)` must add samples [first, first+n) of every pixel into `acc`
    (on the GPU: Context.render into the buffer bound with Context.bind_accum).  Afterwards `acc` on rank `dst` holds all
    `count` samples; on the other ranks it holds partial sums and must be cleared before the next batch."""
    first, n = sample_range(begin, count, rank, world)
    if n:
        render_fn(first, n)
    return reduce_accum(acc, dst=dst, group=group)
