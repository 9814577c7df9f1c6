#!/usr/bin/env python
"""bench.py — headline benchmark: Mrays/s of the path-tracing hot path on BASELINE.json's config C2
(default scene, 1920x1080, 1024 spp, depth 30) on N B200s, beside the reference kernel on the host cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step = one full pass of the hot path over the workload: every pixel x every sample of the job.
ray = one nearest-hit query (path segment), counted by the kernel itself.
N > 1: the job's samples are split across ranks (global sample indices, so the picture does not depend on N),
each rank accumulates a private linear-sum buffer, one NCCL reduce (torch.distributed) sums them on rank 0,
rank 0 finalizes.  `scaling` is therefore "strong".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))

W, H, SPP, DEPTH = 1920, 1080, 1024, 30
TABLE_SEED = 12345
MESH_MODE = 0          # RT_MESH_FIRST_FRONT (the reference's rule); the LBVH configs use RT_MESH_NEAREST
CONFIG = "c2"
SLOTS_PER_RAY = 867.0   # measured FP32 issue slots per ray of the reference algorithm, default scene
SLOTS_SOURCE = ("SURVEY.md 8(d): 1151 unfused FP32 operations per ray - 283 FMA-fusable pairs, default scene; tools/opcount.py "
                "(the reference kernel run with an operation-counting float) re-measures 1152 operations per ray")
METRIC = "Mrays/s at 1080p per scene, 1/2/4/8 B200, vs reference OpenCL on host CPU"
WORKLOAD = "C2: default scene (8 spheres, 1 plane, 1 lens, 2 textured cubes = 24 triangles), 1920x1080, 1024 spp, max depth 30"


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write banners to fd 1 (NCCL prints its version there when the
    box sets NCCL_DEBUG), so fd 1 is pointed at stderr for the whole run and the line goes to the saved descriptor."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp", type=int, default=None)
    ap.add_argument("--arith", default="fast", choices=["fast", "strict"])
    ap.add_argument("--rng", default="counter_table", choices=["counter_table", "counter", "table"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reduce", default="fused", choices=["fused", "nccl"],
                    help="N > 1: fused = every rank reduces+finalizes its rows out of the peers' buffers (CUDA IPC over NVLink, "
                         "rt_finalize_peers); nccl = one torch.distributed reduce to rank 0 + rt_finalize")
    ap.add_argument("--no-heads", action="store_true", help="RT_HEAD_OFF: trace every segment of every sample (A/B of the path-head memoisation)")
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c4", "c5"],
                    help="BASELINE.json configs; the driver's default (and the headline) is c2")
    return ap.parse_args()


def select_config(a):
    """c2 is the bench line; the other BASELINE configs are reported with the same code for DESIGN.md."""
    global W, H, SPP, DEPTH, WORKLOAD, MESH_MODE, CONFIG, SLOTS_PER_RAY, SLOTS_SOURCE
    CONFIG = a.config
    if a.config == "c1":
        W, H, SPP, DEPTH = 800, 600, 16, 8
        WORKLOAD = "C1: default scene, 800x600, 16 spp, max depth 8"
    elif a.config == "c3":
        W, H, SPP, DEPTH = 1920, 1080, 4096, 30
        WORKLOAD = "C3: reference cube meshes (textured) + diffuse/dielectric/reflective spheres, 1920x1080, 4096 spp, max depth 30"
        # tools/opcount.py c3: 1005 operations per ray, 224 fusable pairs by its own (stricter) pairing rule; scaled to the
        # SURVEY's pairing convention (283 pairs where the tool finds 237 on the default scene): 1005 - 224 * 283 / 237
        SLOTS_PER_RAY = 738.0
        SLOTS_SOURCE = ("tools/opcount.py c3 (the reference kernel run with an operation-counting float): 1005 FP32 operations per ray, "
                        "224 FMA-fusable pairs scaled by 283/237 to SURVEY.md 8(d)'s pairing convention")
    elif a.config == "c4":
        W, H, SPP, DEPTH, MESH_MODE = 3840, 2160, 64, 30, 1
        WORKLOAD = "C4: synthetic closed mesh of 1 310 720 triangles + 1001 spheres + floor plane, 3840x2160, GPU-built LBVH, max depth 30"
        SLOTS_PER_RAY = None
    elif a.config == "c5":
        W, H, SPP, DEPTH, MESH_MODE = 3840, 2160, 64, 16, 1
        WORKLOAD = "C5: 100 001 mixed-material spheres + floor plane, 3840x2160, GPU-built LBVH, max depth 16"
        SLOTS_PER_RAY = None
    if a.spp is None:
        a.spp = SPP


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def load_inputs():
    import rtb200
    from rtb200 import fixtures as fx
    if CONFIG in ("c4", "c5"):
        import numpy as np
        from rtb200 import synth
        arr, _ = synth.c4_scene(subdiv=8) if CONFIG == "c4" else synth.c5_scene()
        tex = np.zeros((1, 1, 1, 4), np.float32)
    else:
        arr, tex8 = fx.c3_scene() if CONFIG == "c3" else fx.default_scene()
        tex = fx.textures_to_float(tex8)
    cam = fx.camera(60, W / H)
    table = rtb200.make_random_table(TABLE_SEED)
    return arr, tex, cam, table


def cpu_reference_rate(spp_per_step, steps, warmup, first=0):
    """The reference's own kernel text compiled as C++ (oracle/_ref/libref_fast.so), all host threads, one
    NDRange-equivalent pass per sample like render()/renderAgain().  Returns (Mrays/s, cores, kind, sample)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle_py as op
    arr, tex, cam, table = load_inputs()
    cores = os.cpu_count() or 1
    try:
        R, kind = op.RefOracle("fast"), "reference"
        def run(first, n, img):
            _, rays, secs = R.render(d, tex, cam, table, W, H, first, n, depth=DEPTH, threads=cores, image=img)
            return rays, secs
    except (FileNotFoundError, OSError):
        P, kind = op.PortOracle(), "port"
        def run(first, n, img):
            _, rays, secs = P.render(d, tex, cam, table, W, H, first, n, P.params(max_depth=DEPTH, threads=cores), image=img)
            return rays, secs
    d = op.make_desc(arr)
    img = np.zeros((H, W, 4), np.float32)
    k = first
    for _ in range(warmup):
        run(k, spp_per_step, img); k += spp_per_step
    rays_t, secs_t, per_step = 0, 0.0, []
    for _ in range(steps):
        r, s = run(k, spp_per_step, img); k += spp_per_step
        rays_t += r; secs_t += s; per_step.append(s)
    return rays_t / secs_t / 1e6, cores, kind, rays_t, secs_t, per_step


def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    spp_step = 8      # a bounded sample of the workload per step: ~0.6 s of all host cores on C2
    rate, cores, kind, rays, secs, per = cpu_reference_rate(spp_step, a.steps, a.warmup)
    sample = "%dx%d, %d spp per step (one pass per sample), depth %d, %d host threads" % (W, H, spp_step, DEPTH, cores)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(rate, 3), "unit": "Mrays/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(1e3 * secs / max(a.steps, 1), 3),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": round(rate, 3), "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(rate, 3), "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), file=JSON_OUT, flush=True)


def main_b200(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    import rtb200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    arith = rtb200.ARITH_FAST if a.arith == "fast" else rtb200.ARITH_STRICT
    rng = {"counter": rtb200.RNG_COUNTER, "table": rtb200.RNG_TABLE, "counter_table": rtb200.RNG_COUNTER_TABLE}[a.rng]

    arr, tex, cam, table = load_inputs()
    ctx = rtb200.Context(W, H, local)
    ctx.set_params(max_depth=DEPTH, rng_mode=rng, seed=1, arith=arith, mesh_mode=MESH_MODE,
                   path_head=rtb200.HEAD_OFF if a.no_heads else rtb200.HEAD_ON)
    ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    fused = world > 1 and a.reduce == "fused"
    accum = None
    if not fused:          # torch-owned buffer (what dist.reduce needs); the fused path exports the context's own buffer
        accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        ctx.set_accum_device_buffer(accum.data_ptr(), accum.numel() * 4)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    from rtb200.multigpu import PeerFinalize, reduce_to_rank0, sample_range
    peers = PeerFinalize(ctx, stream_ordered=True) if fused else None
    spp = a.spp
    first, n_mine = sample_range(spp, world, rank)

    peak_gslots, nominal_mhz = rtb200.measure_fp32_peak(local)

    queries_steps = []

    def step():
        if fused:
            ctx.clear_accum()
        else:
            accum.zero_()
        ctx.invalidate_caches()          # every step rebuilds the path heads: nothing is carried over between steps
        ctx.render_samples(cam, first, n_mine, rtb200.ACCUM_SUM)
        st = ctx.stats()
        rays, kms = st.rays_last, st.ms_trace_last
        queries_steps.append(st.queries_last)
        if fused:
            peers.finalize(spp)
        else:
            reduce_to_rank0(accum)
            if rank == 0:
                ctx.finalize(spp)
        return rays, kms

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        step(); flush.zero_()
    launches0 = ctx.stats().kernel_launches
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    rays_steps, kernel_ms = [], []
    for _ in range(a.steps):
        r, kms = step()
        rays_steps.append(r); kernel_ms.append(kms)
        flush.zero_()
    e1.record(stream)
    sync_all()
    elapsed_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.stats().kernel_launches - launches0

    t = torch.tensor([elapsed_ms, float(sum(rays_steps)), float(launches)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        elapsed_ms, rays_total, launches = tmax[0].item(), tsum[1].item(), int(tsum[2].item())
    else:
        rays_total = t[1].item()
    value = rays_total / (elapsed_ms * 1e-3) / 1e6

    # ---- end to end through the C ABI with HOST buffers: every step uploads scene, textures, table (pinned
    # host memory -> device), renders, reduces, finalizes and reads the image back to the host.
    img_pinned = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()   # the caller's host image buffer (pinned, like the inputs)
    img_host = img_pinned.numpy()
    tex_pinned = torch.from_numpy(tex).pin_memory()
    table_pinned = torch.from_numpy(table).pin_memory()
    h2d = tex_pinned.numel() * 4 + table_pinned.numel() * 4 + sum(v.nbytes for v in arr.values()) + 48
    d2h = img_host.nbytes if rank == 0 else 0

    def e2e_step():
        ctx.upload_textures(tex_pinned.numpy()); ctx.upload_scene(arr); ctx.set_random_table(table_pinned.numpy())
        r, _ = step()
        if rank == 0:
            ctx.read_image(img_host)
        return r

    e2e_step()
    sync_all()
    t0 = time.perf_counter()
    e2e_rays = 0
    e2e_steps = max(1, min(a.steps, 3))
    for _ in range(e2e_steps):
        e2e_rays += e2e_step()
    sync_all()
    e2e_s = time.perf_counter() - t0
    t2 = torch.tensor([e2e_s, float(e2e_rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        m = t2.clone(); dist.all_reduce(m, op=dist.ReduceOp.MAX)
        s = t2.clone(); dist.all_reduce(s, op=dist.ReduceOp.SUM)
        e2e_s, e2e_rays = m[0].item(), s[1].item()
    e2e_value = e2e_rays / e2e_s / 1e6

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        k_ms = sum(kernel_ms) / len(kernel_ms)
        rays_launch = rays_steps[0]
        achieved = (SLOTS_PER_RAY or 0.0) * rays_launch / (k_ms * 1e-3) / 1e9
        queries_launch = queries_steps[-1]
        executed = (SLOTS_PER_RAY or 0.0) * queries_launch / (k_ms * 1e-3) / 1e9
        bvh = int(ctx.stats().bvh_nodes) > 0
        kernel = ("k_trace_%s_bvh" if bvh else ("k_trace_%s" if a.no_heads else "k_trace_tail_%s")) % a.arith
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        alg_bytes = 2.0 * W * H * 16          # accumulation buffer read + write, once per launch
        out = {
            "metric": METRIC, "value": round(value, 1), "unit": "Mrays/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(elapsed_ms / a.steps, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "spp_total": spp, "spp_per_gpu": n_mine, "arith": a.arith, "rng": a.rng,
                       "parallelism": ("spp-split x%d + %s" % (world, "fused peer-memory reduce+finalize (rt_finalize_peers)" if fused else "1 NCCL reduce"))
                                      if world > 1 else "single GPU",
                       "l2": "flushed between steps (256 MiB memset); inputs are < 40 MB, the kernel is issue bound",
                       "rays_per_step": int(rays_total / a.steps),
                       "ray": "one path segment = one hitScene call of the reference (kernels/raytracer.cl:322-360); the image is the "
                              "reference's, sample for sample",
                       "path_heads": (not a.no_heads) and not bvh,
                       "queries_executed_per_step_rank0": int(queries_launch)},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                    "what": "rt_upload_textures + rt_upload_scene + rt_set_random_table from pinned host memory, "
                            "rt_render_samples, reduce, rt_finalize, rt_read_image to host, every step"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32_issue", "achieved": round(achieved, 1), "peak": round(peak_gslots, 1),
                         "unit": "G lane-slots/s", "frac": round(achieved / peak_gslots, 4) if (peak_gslots and SLOTS_PER_RAY) else None,
                         "traffic": (161039360 if (CONFIG == "c2" and spp == 1024 and world == 1 and not a.no_heads and a.arith == "fast") else None),
                         "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of one k_trace_tail_fast launch of this command, bytes, from "
                                         "profiles/r01_v7b_k_trace_tail_fast_ncu_summary.txt (ncu --set full, the final round-1 build); null for other configurations",
                         "kernel": kernel, "kernel_ms": round(k_ms, 3),
                         "achieved_executed": round(executed, 1),
                         "frac_executed": round(executed / peak_gslots, 4) if (peak_gslots and SLOTS_PER_RAY) else None,
                         "note": "achieved = slots_per_ray x path segments of the launch (the reference algorithm's work for this image) / "
                                 "kernel time; achieved_executed counts only the nearest-hit queries the kernel ran: the head of a "
                                 "path (same for every sample of a pixel) is traced once per pixel, not once per sample, so frac can exceed 1",
                         "slots_per_ray": SLOTS_PER_RAY, "slots_source": SLOTS_SOURCE if SLOTS_PER_RAY else None,
                         "rays_per_launch": int(rays_launch),
                         "peak_source": "measured live: dependent-free FFMA loop on all SMs (rt_measure_fp32_peak); "
                                        "MEASURED_PEAKS.json has no FP32 entry",
                         "hbm": {"achieved_gbs": round(alg_bytes / (k_ms * 1e-3) / 1e9, 2), "peak_gbs": hbm_peak,
                                 "note": "evidence only: 33 MB per launch, not the bound"}},
        }
        out["config"]["bvh_nodes"] = int(ctx.stats().bvh_nodes)
        if world == 1 and not a.no_cpu_baseline and CONFIG in ("c1", "c2", "c3"):
            rate, cores, kind, rays, secs, _ = cpu_reference_rate(32, 4, 1)
            out["cpu_baseline"] = {"value": round(rate, 3), "unit": "Mrays/s", "cores": cores, "kind": kind,
                                   "sample": "%dx%d, 128 timed spp in 4 steps (one pass per sample), depth %d, %.1f s of CPU work"
                                             % (W, H, DEPTH, secs)}
        print(json.dumps(out), file=JSON_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    JSON_OUT = claim_stdout()
    select_config(args)
    if args.impl == "reference":
        main_reference(args)
    else:
        main_b200(args)
