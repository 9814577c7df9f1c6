"""The REAL multi-process path of the sample split: one process per GPU (torchrun, NCCL), CUDA IPC handles exchanged with
rt_ipc_export / rt_ipc_open_peers, every rank renders its global sample range, rt_finalize_peers_async sums the ranks'
buffers over NVLink peer loads into rank 0's image.  The N-rank picture must equal the picture rank 0 renders alone to
float-summation tolerance (2e-6 on the sqrt(mean) image) — for the fused exchange (between host barriers, between NCCL
barriers on the stream, and with the device-side arrival-flag barrier) and for the NCCL reduce.
Skipped on boxes with one GPU (bench.py runs the same check at every N > 1 and prints it as "multi_gpu_check")."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _main(out_path):
    sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
    import torch
    import torch.distributed as dist
    import rtb200
    from rtb200 import fixtures as fx
    from rtb200.multigpu import PeerFinalize, reduce_to_rank0, split_image_check
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    torch.cuda.set_stream(torch.cuda.Stream())       # a real stream: the null handle of torch's default stream would mean "private stream"
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    arr, tex8 = fx.default_scene()
    tex = fx.textures_to_float(tex8)
    W, H = 640, 360
    cam = fx.camera(60, W / H)
    table = rtb200.make_random_table(12345)
    results = {}
    for arith in ("strict", "fast"):
        for mode in ("fused", "fused_stream", "fused_flags", "nccl"):
            ctx = rtb200.Context(W, H, local)
            ctx.set_params(max_depth=30, rng_mode=rtb200.RNG_COUNTER_TABLE, seed=1,
                           arith=rtb200.ARITH_STRICT if arith == "strict" else rtb200.ARITH_FAST)
            ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
            if mode == "nccl":
                accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
                ctx.set_accum_device_buffer(accum.data_ptr(), accum.numel() * 4)

                def fin(spp):
                    torch.cuda.synchronize()
                    reduce_to_rank0(accum)
                    torch.cuda.synchronize()
                    if rank == 0:
                        ctx.finalize(spp)
            else:
                if mode != "fused":
                    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
                peers = PeerFinalize(ctx, stream_ordered=(mode == "fused_stream"), device_barrier=(mode == "fused_flags"))
                fin = peers.finalize
            r = split_image_check(ctx, cam, 24, world, rank, fin)
            rp = split_image_check(ctx, cam, 24, world, rank, fin, split="pixels")   # interleaved pixel groups: identical picture
            ctx.set_pixel_split(1, 0)
            if mode == "fused_flags":
                for _ in range(3):                      # the epochs keep counting: repeated rounds on the same contexts
                    r2 = split_image_check(ctx, cam, 24, world, rank, fin)
                    if rank == 0:
                        r["ok"] = r["ok"] and r2["ok"]
                timed_out = ctx.peer_barrier_timed_out()
                if rank == 0:
                    r["ok"] = r["ok"] and not timed_out
            if rank == 0:
                results["%s/%s" % (arith, mode)] = r
                results["%s/%s/pixels" % (arith, mode)] = rp
            if mode != "nccl":
                peers.close()
            ctx.close()
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump(results, f)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_rank_ipc_image_equals_single_rank_image(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (bench.py prints the same check as multi_gpu_check at every N > 1)")
    out = str(tmp_path / "mp.json")
    port = 29600 + os.getpid() % 2000
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__), out],
                          timeout=600)
    res = json.load(open(out))
    print(res)
    assert len(res) == 16
    for key, r in res.items():
        assert r["ok"], (key, r)
        if key.endswith("/pixels") and key.startswith("strict/"):
            assert r["max_abs"] == 0.0, (key, r)


if __name__ == "__main__":
    _main(sys.argv[1])
