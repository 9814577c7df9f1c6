"""rtb200.multigpu — the sample split across the GPUs of one box (one process per GPU).

The path shards by independent units: rank r of `world` renders the global sample indices
[r*spp/world, (r+1)*spp/world) of the whole frame into a private linear-sum buffer (RT_ACCUM_SUM); the only
exchange step is ONE sum-reduce of the W*H float4 buffers to rank 0, which then applies the reference's
tail (divide by spp, sqrt) with rt_finalize.  Because sample indices are global and the counter RNG is keyed by
(pixel, sample, bounce), the picture does not depend on `world` beyond float summation order."""
from __future__ import annotations


def sample_range(spp: int, world: int, rank: int):
    """-> (first, count): contiguous, disjoint, covering [0, spp); counts differ by at most one."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    first = rank * spp // world
    return first, (rank + 1) * spp // world - first


def reduce_to_rank0(accum, group=None):
    """Sum the per-rank accumulation buffers onto rank 0 (NCCL over NVLink on GPUs; gloo in the CPU tests)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM, group=group)
    return accum
