"""rtb200.multigpu — the sample split across the GPUs of one box (one process per GPU).

The path shards by independent units: rank r of `world` renders the global sample indices
[r*spp/world, (r+1)*spp/world) of the whole frame into a private linear-sum buffer (RT_ACCUM_SUM); the only
exchange step is ONE sum-reduce of the W*H float4 buffers to rank 0, which then applies the reference's
tail (divide by spp, sqrt) with rt_finalize.  Because sample indices are global and the counter RNG is keyed by
(pixel, sample, bounce), the picture does not depend on `world` beyond float summation order."""
from __future__ import annotations

# Two exchange steps are provided:
#   reduce_to_rank0      one NCCL sum-reduce of the buffers (torch.distributed), rt_finalize on rank 0;
#   PeerFinalize         fused: every rank reduces + finalizes its slice of rows straight out of the peers'
#                        buffers (CUDA IPC over NVLink) into rank 0's image — rt_finalize_peers.


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


class PeerFinalize:
    """Fused reduce + finalize over peer memory.  Handles are exchanged once (all_gather_object); per render every
    rank calls finalize(spp) between two barriers."""

    def __init__(self, ctx, group=None, stream_ordered=False, device_barrier=False):
        """stream_ordered: the context renders on torch's current CUDA stream (rt_set_stream); the two rank barriers
        are then tiny all-reduces ordered on that stream and nothing blocks the host until the caller synchronises.
        device_barrier: no collective at all — the ranks meet through arrival words in peer memory
        (rt_finalize_peers_barrier_async); everything stays on the context's stream."""
        import torch.distributed as dist
        self.ctx, self.group, self.stream_ordered, self.device_barrier = ctx, group, stream_ordered, device_barrier
        self.flag = None
        if stream_ordered:
            import torch
            self.flag = torch.zeros(1, dtype=torch.float32, device="cuda")
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        mine = ctx.ipc_export()
        allh = [None] * self.world
        dist.all_gather_object(allh, mine, group=group)
        ctx.ipc_open_peers(self.rank, [h[0] for h in allh], allh[0][1])

    def finalize(self, spp):
        import torch.distributed as dist
        if self.device_barrier:
            self.ctx.finalize_peers_barrier(spp)
            return
        if self.stream_ordered:
            dist.all_reduce(self.flag, group=self.group)      # every rank's render kernel is behind this on its stream
            self.ctx.finalize_peers(spp, sync=False)
            dist.all_reduce(self.flag, group=self.group)      # every slice of rank 0's image is behind this
            return
        dist.barrier(group=self.group)       # every rank's render has finished (render calls are synchronous)
        self.ctx.finalize_peers(spp)
        dist.barrier(group=self.group)       # every slice of rank 0's image has been written

    def close(self):
        self.ctx.close_peers()


def split_image_check(ctx, cam, spp, world, rank, finalize_split, group=None):
    """Correctness of the real multi-process path: the `world`-rank picture (every rank renders its global sample range
    into its linear-sum buffer, finalize_split(spp) = the exchange step: PeerFinalize.finalize or reduce_to_rank0 +
    rt_finalize on rank 0) against the same job rendered by rank 0 alone.  The RNG is keyed by global sample indices, so
    the two differ by float summation order only.  Collective: every rank calls it; rank 0 gets
    {"max_abs", "spp", "ok"}, the others None."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from . import ACCUM_SUM
    first, n = sample_range(spp, world, rank)
    ctx.clear_accum()
    ctx.render_samples(cam, first, n, ACCUM_SUM)
    finalize_split(spp)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier(group=group)
    out = None
    if rank == 0:
        split = ctx.read_image().copy()
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, ACCUM_SUM)
        ctx.finalize(spp)
        single = ctx.read_image()
        max_abs = float(np.abs(split[..., :3] - single[..., :3]).max())
        out = {"max_abs": max_abs, "spp": int(spp), "tolerance": 2e-6,
               "ok": bool(max_abs <= 2e-6 and np.all(split[..., 3] == 1.0) and float(single[..., :3].max()) > 0.1)}
    if world > 1:
        dist.barrier(group=group)
    return out
