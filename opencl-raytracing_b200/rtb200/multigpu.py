"""rtb200.multigpu — splitting one render job across the GPUs of one box (one process per GPU).

Two splits of the independent units (pixel-samples):
  "spp"     (north_star) rank r renders the global sample indices [r*spp/world, (r+1)*spp/world) of the WHOLE frame;
  "pixels"  rank r renders ALL samples of the interleaved 128-pixel groups g with g % world == r (rt_set_pixel_split):
            per-lane work items stay as long as on one GPU (the kernel's efficiency depends on that), nothing is computed
            twice, and the summed picture is bit-identical to one GPU's in the strict arm.  Measured SLOWER than the spp
            split on C2 (2 GPUs: 128 vs 136 Grays/s, 8 GPUs: 417 vs 484): not the default.
Either way every rank holds a linear-sum buffer and ONE exchange step produces the image on rank 0.


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
        if stream_ordered and not device_barrier:
            import torch
            # rt_set_stream(NULL) means "the context's private stream", and torch's DEFAULT stream is the null handle: a context
            # given that handle renders on its own stream and nothing orders its kernels with the collectives (found by
            # tests/test_gpu_multiprocess.py: rank 0 finalized before rank 1 had rendered).  Demand a real stream.
            if torch.cuda.current_stream().cuda_stream == 0:
                raise ValueError("PeerFinalize(stream_ordered=True) needs a non-default current torch stream "
                                 "(torch.cuda.set_stream(torch.cuda.Stream())) shared with the context through rt_set_stream")
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


def split_image_check(ctx, cam, spp, world, rank, finalize_split, group=None, split="spp"):
    """Correctness of the real multi-process path: the `world`-rank picture (every rank renders its global sample range
    into its linear-sum buffer, finalize_split(spp) = the exchange step: PeerFinalize.finalize or reduce_to_rank0 +
    rt_finalize on rank 0) against the same job rendered by rank 0 alone.  The RNG is keyed by global sample indices, so
    the two differ by float summation order only.  split = "pixels": the ranks render interleaved pixel groups
    (rt_set_pixel_split) with all samples each — every pixel then has exactly one contribution and the pictures are
    IDENTICAL in the strict arm (the fast arm may plan its sample chunks differently for a part of the frame).  Collective: every rank calls it; rank 0 gets {"max_abs", "spp", "ok"}, the others None."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from . import ACCUM_SUM
    first, n = sample_range(spp, world, rank) if split == "spp" else (0, spp)
    if split == "pixels":
        ctx.set_pixel_split(world, rank)
    ctx.clear_accum()
    ctx.render_samples(cam, first, n, ACCUM_SUM)
    finalize_split(spp)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier(group=group)
    out = None
    if rank == 0:
        split_img = ctx.read_image().copy()
        ctx.set_pixel_split(1, 0)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, ACCUM_SUM)
        ctx.finalize(spp)
        single = ctx.read_image()
        max_abs = float(np.abs(split_img[..., :3] - single[..., :3]).max())
        # pixel split: every pixel has ONE contribution, so the pictures are identical whenever both runs add a pixel's samples
        # in the same order (strict arm; fast arm with the same sample-chunk plan) — else they differ by summation order, too
        tol = 2e-6
        out = {"max_abs": max_abs, "spp": int(spp), "tolerance": tol, "split": split, "bit_identical": bool(max_abs == 0.0),
               "ok": bool(max_abs <= tol and np.all(split_img[..., 3] == 1.0) and float(single[..., :3].max()) > 0.1)}
        if split == "pixels":
            ctx.set_pixel_split(world, rank)
    if world > 1:
        dist.barrier(group=group)
    return out


class SharedHostImage:
    """One host image (H x W x 4 float32) that EVERY rank's finalize kernel can store into: POSIX shared memory created by
    rank 0, attached by the others, page-locked and mapped into each rank's device address space (cudaHostRegister).  With
    rt_set_finalize_target every rank writes its rows of the picture straight to the host over its own PCIe link — the
    read-back of a frame uses all links of the box instead of rank 0's alone.  world == 1: ordinary pinned memory."""

    def __init__(self, H, W, world=1, rank=0, group=None):
        import numpy as np
        import torch
        self.shm, self.registered, self.nbytes = None, False, H * W * 16
        if world == 1:
            self.tensor = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
            self.array = self.tensor.numpy()
            self.ptr = self.tensor.data_ptr()
            return
        import torch.distributed as dist
        from multiprocessing import shared_memory
        name = [None]
        if rank == 0:
            self.shm = shared_memory.SharedMemory(create=True, size=self.nbytes)
            name[0] = self.shm.name
        dist.broadcast_object_list(name, src=0, group=group)
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=name[0])
            try:        # Python < 3.13 registers an ATTACHED segment with this process' resource tracker, which then tries to
                from multiprocessing import resource_tracker          # unlink it at exit; the segment belongs to rank 0
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.array = np.ndarray((H, W, 4), dtype=np.float32, buffer=self.shm.buf)
        self.ptr = self.array.ctypes.data
        err = torch.cuda.cudart().cudaHostRegister(self.ptr, self.nbytes, 3)      # cudaHostRegisterPortable | cudaHostRegisterMapped
        if int(err) != 0:
            raise RuntimeError("cudaHostRegister of the shared host image failed: %s" % (err,))
        self.registered = True
        self.rank = rank
        dist.barrier(group=group)

    def close(self):
        import torch
        if self.registered:
            torch.cuda.cudart().cudaHostUnregister(self.ptr)
            self.registered = False
        if self.shm is not None:
            self.array = None
            try:
                self.shm.close()
                if self.rank == 0:
                    self.shm.unlink()
            except Exception:
                pass
            self.shm = None
