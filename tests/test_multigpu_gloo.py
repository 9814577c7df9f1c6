"""CPU test of the N > 1 path: world_size 2 over gloo.  The renderer on each rank is the oracle (there is no GPU
here); what is under test is the product's host logic — rtb200.multigpu.sample_range and reduce_to_rank0 — and
the property the split relies on: disjoint global sample ranges sum to the single-process result."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sample_range_partitions():
    from rtb200.multigpu import sample_range
    for spp in (1, 7, 16, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            parts = [sample_range(spp, world, r) for r in range(world)]
            assert parts[0][0] == 0 and sum(n for _, n in parts) == spp
            for (f0, n0), (f1, _) in zip(parts, parts[1:]):
                assert f0 + n0 == f1
            assert max(n for _, n in parts) - min(n for _, n in parts) <= 1
    with pytest.raises(ValueError):
        sample_range(8, 2, 2)


def _worker(rank, world, port, out_path):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
    import torch
    import torch.distributed as dist
    import oracle_py as op
    from rtb200 import fixtures as fx
    from rtb200.multigpu import reduce_to_rank0, sample_range
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    arr, tex8 = fx.default_scene()
    tex = fx.textures_to_float(tex8)
    W, H, spp = 64, 48, 9
    cam = fx.camera(60, W / H)
    P = op.PortOracle()
    prm = P.params(max_depth=30, rng_mode=op.RNG_COUNTER, seed=3, accum_mode=op.ACCUM_SUM, threads=2)
    first, n = sample_range(spp, world, rank)
    mine, _, _ = P.render(op.make_desc(arr), tex, cam, None, W, H, first, n, prm)
    t = torch.from_numpy(mine.copy())
    reduce_to_rank0(t)
    if rank == 0:
        whole, _, _ = P.render(op.make_desc(arr), tex, cam, None, W, H, 0, spp, prm)
        np.savez(out_path, reduced=t.numpy(), whole=whole)
    dist.barrier()
    dist.destroy_process_group()


def test_split_plus_reduce_equals_single(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r.npz")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    z = np.load(out)
    assert np.all(z["whole"][..., 3] == 9) and np.all(z["reduced"][..., 3] == 9)
    assert np.abs(z["reduced"] - z["whole"]).max() <= 1e-5
