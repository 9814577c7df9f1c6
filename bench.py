#!/usr/bin/env python
"""bench.py — headline benchmark: Mrays/s of the path-tracing hot path on BASELINE.json's config C2
(default scene, 1920x1080, 1024 spp, depth 30) on N B200s, beside the reference kernel on the host cores.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config c1..c5] [--no-extras]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step = one full pass of the hot path over the workload: every pixel x every sample of the job.
Two rates are printed and must not be confused:
  value           reference-equivalent Mrays/s: path segments of the reference algorithm for this image (one hitScene call
                  of kernels/raytracer.cl:322-360 each) per second — the time-to-same-image rate against the reference arm;
  value_executed  nearest-hit queries the device actually ran per second (the head of a pixel's path is the same for every
                  sample and is traced once per pixel, so fewer queries run than the reference makes).
roofline.frac is the EXECUTED work over the measured FP32-issue peak; roofline.algorithmic_gain = segments / queries.
N > 1: the job's samples are split across ranks (global sample indices, so the picture does not depend on N),
each rank accumulates a private linear-sum buffer, the exchange step sums them into rank 0's image (fused peer-memory
kernel, or one NCCL reduce) — `scaling` is "strong"; every N > 1 run checks the N-rank picture against rank 0 alone.
The default run (config c2) also measures the other BASELINE configs briefly and reports them under "extra_configs".
"""
import argparse
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "opencl-raytracing_b200")

TABLE_SEED = 12345
METRIC = "Mrays/s at 1080p per scene, 1/2/4/8 B200, vs reference OpenCL on host CPU"
RAY = ("value = reference-equivalent rate: one ray = one path segment = one hitScene call of the reference "
       "(kernels/raytracer.cl:322-360) for this image, sample for sample; value_executed = nearest-hit queries the device ran "
       "(path heads are traced once per pixel, not once per sample)")
# FP32 issue slots per primitive test / per ray of the reference algorithm, SURVEY.md 8(d)
SLOT_SPHERE, SLOT_PLANE, SLOT_LENS, SLOT_TRI, SLOT_BOX, SLOT_RAY = 12.0, 17.0, 24.0, 30.0, 16.0, 60.0

CONFIGS = {
    "c1": dict(W=800, H=600, spp=16, depth=8, mesh_mode=0, scene="default", slots=867.0, short_spp=16,
               workload="C1: default scene, 800x600, 16 spp, max depth 8 (BASELINE.json configs[0], the reference's CPU-runnable case)"),
    "c2": dict(W=1920, H=1080, spp=1024, depth=30, mesh_mode=0, scene="default", slots=867.0, short_spp=1024,
               workload="C2: default scene (8 spheres, 1 plane, 1 lens, 2 textured cubes = 24 triangles), 1920x1080, 1024 spp, max depth 30"),
    "c3": dict(W=1920, H=1080, spp=4096, depth=30, mesh_mode=0, scene="c3", slots=738.0, short_spp=1024,
               workload="C3: reference cube meshes (textured) + diffuse/dielectric/reflective spheres, 1920x1080, 4096 spp, max depth 30"),
    "c4": dict(W=3840, H=2160, spp=64, depth=30, mesh_mode=1, scene="c4", slots=None, short_spp=256,
               workload="C4: synthetic closed mesh of 1 310 720 triangles + 1001 spheres + floor plane, 3840x2160, GPU-built LBVH, max depth 30"),
    "c5": dict(W=3840, H=2160, spp=64, depth=16, mesh_mode=1, scene="c5", slots=None, short_spp=256,
               workload="C5: 100 001 mixed-material spheres + floor plane, 3840x2160, GPU-built LBVH, max depth 16"),
}
SLOTS_SOURCE = {
    867.0: "SURVEY.md 8(d): 1151 unfused FP32 operations per ray - 283 FMA-fusable pairs, default scene; tools/opcount.py "
           "(the reference kernel run with an operation-counting float) re-measures 1152 operations per ray",
    738.0: "tools/opcount.py c3 (the reference kernel run with an operation-counting float): 1005 FP32 operations per ray, "
           "224 FMA-fusable pairs scaled by 283/237 to SURVEY.md 8(d)'s pairing convention",
    None: "COUNTED work of the run (SURVEY.md 8d, BVH rule): 16 x box tests + 12 x sphere + 24 x lens + 30 x triangle + 17 x plane "
          "leaf/list tests + 60 x queries, counters from the instrumented untimed copy of the render kernel (rt_get_work_counts)",
}
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel from the ncu --set full captures under
# profiles/ (r02_*_ncu_summary.txt), keyed by (config, spp per launch); null where no capture exists for that shape.
NCU_TRAFFIC = {}
try:
    NCU_TRAFFIC = {tuple(k.split(":")): v for k, v in json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json"))).items()}
except Exception:
    pass


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
    ap.add_argument("--reduce", default="fused", choices=["fused", "fused_nccl_barriers", "nccl"],
                    help="N > 1: fused = every rank reduces+finalizes its rows out of the peers' buffers (CUDA IPC over NVLink, "
                         "rt_finalize_peers_barrier_async: the ranks meet through arrival words in peer memory, no collective call); "
                         "fused_nccl_barriers = the same kernel between two NCCL all-reduces of a scalar (round 1); "
                         "nccl = one torch.distributed reduce to rank 0 + rt_finalize")
    ap.add_argument("--split", default="spp", choices=["pixels", "spp"],
                    help="N > 1: pixels = every rank renders all samples of its interleaved 128-pixel groups (rt_set_pixel_split; picture bit-identical "
                         "in the strict arm, equal to rounding in the fast arm; work items as long as on one GPU); spp = every rank renders its share of the samples of the whole frame "
                         "(north_star's split, round 1)")
    ap.add_argument("--host-image", default="shared", choices=["shared", "rank0"],
                    help="end-to-end read-back: rank0 = the exchange step writes rank 0's device image, rank 0 copies it to its pinned host "
                         "buffer (rt_read_image); shared = every rank's finalize kernel stores its rows straight into ONE page-locked host image "
                         "in shared memory over its own PCIe link (rt_set_finalize_target)")
    ap.add_argument("--no-heads", action="store_true", help="RT_HEAD_OFF: trace every segment of every sample (A/B of the path-head memoisation)")
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS),
                    help="BASELINE.json configs; the driver's default (and the headline) is c2")
    ap.add_argument("--no-extras", action="store_true", help="skip the brief runs of the other BASELINE configs (extra_configs)")
    return ap.parse_args()


class ClockSampler:
    """SM clock, power and throttle reasons DURING the timed region (B200_PROFILING.md's clocks line).  NVML is polled in-process
    every 5 ms — at 8 GPUs a timed region is 50 ms, which `nvidia-smi -lms 100` misses altogether (0 samples); nvidia-smi remains
    the fallback where the NVML binding is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index
        self.nv, self.h, self.stop_flag, self.samples = None, None, False, []

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # LOCAL_RANK indexes the visible devices; NVML wants the physical one
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.gpu]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else self.gpu
            self.h, self.nv = pynvml.nvmlDeviceGetHandleByIndex(idx), pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _poll(self):
        nv, h = self.nv, self.h
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
                pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((sm, mx, pw, int(rs)))
            except Exception:
                pass
            time.sleep(0.005)

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.nv is not None:
            self.stop_flag = True
            self.t.join(timeout=2)
            nv, sm = self.nv, sorted(x[0] for x in self.samples)
            bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
            for name, attr in (("hw_slowdown", "nvmlClocksEventReasonHwSlowdown"), ("sw_power_cap", "nvmlClocksEventReasonSwPowerCap"),
                               ("sw_thermal_slowdown", "nvmlClocksEventReasonSwThermalSlowdown"),
                               ("hw_thermal_slowdown", "nvmlClocksEventReasonHwThermalSlowdown")):
                bits[name] = int(getattr(nv, attr, bits[name]))
            reasons = sorted(n for n, b in bits.items() if any(x[3] & b for x in self.samples))
            return {"sm_mhz": float(sm[len(sm) // 2]) if sm else None, "sm_max_mhz": float(max(x[1] for x in self.samples)) if sm else None,
                    "power_w_max": max(x[2] for x in self.samples) if sm else None, "samples": len(sm), "reasons": reasons,
                    "source": "NVML, 5 ms period"}
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
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons), "source": "nvidia-smi -lms 20"}


# ------------------------------------------------------------------------------------------------ inputs
def _module_by_path(name, path):
    """Load a pure-numpy helper (scene fixtures / synthetic scenes) WITHOUT importing the rtb200 package: the reference arm
    must not map the product's shared library, not even for input preparation."""
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


_FX = _SYNTH = None


def helpers():
    global _FX, _SYNTH
    if _FX is None:
        _FX = _module_by_path("_bench_fixtures", os.path.join(PKG, "rtb200", "fixtures.py"))
        _SYNTH = _module_by_path("_bench_synth", os.path.join(PKG, "rtb200", "synth.py"))
    return _FX, _SYNTH


_SCENES = {}


def load_scene(cfg):
    """-> (arrays in the reference's host layouts, float textures [L, H, W, 4], camera 12 floats).  numpy only."""
    import numpy as np
    fx, synth = helpers()
    key = cfg["scene"]
    if key not in _SCENES:
        if key in ("c4", "c5"):
            arr, _ = synth.c4_scene(subdiv=8) if key == "c4" else synth.c5_scene()
            tex = np.zeros((1, 1, 1, 4), np.float32)
        else:
            arr, tex8 = fx.c3_scene() if key == "c3" else fx.default_scene()
            tex = fx.textures_to_float(tex8)
        _SCENES[key] = (arr, tex)
    arr, tex = _SCENES[key]
    return arr, tex, fx.camera(60, cfg["W"] / cfg["H"])


# ------------------------------------------------------------------------------------------------ reference arm
def cpu_reference_rate(cfg, spp_per_step, steps, warmup, window=None):
    """The reference's own kernel text compiled as C++ (oracle/_ref/libref_fast.so), all host threads, one
    NDRange-equivalent pass per sample like render()/renderAgain().  LBVH configs (C4, C5) have no reference-kernel run —
    the reference's first-front-facing mesh rule cannot be answered by a hierarchy and its frame is the full NDRange — so
    they time the oracle port's brute force with the nearest-front-facing rule on a centred window.
    -> dict(rate Mrays/s, cores, kind, rays, secs, sample)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle_py as op
    arr, tex, cam = load_scene(cfg)
    W, H, depth = cfg["W"], cfg["H"], cfg["depth"]
    cores = os.cpu_count() or 1
    d = op.make_desc(arr)
    if cfg["mesh_mode"] == 1:
        P, kind = op.PortOracle(), "port"
        table = P.make_table(TABLE_SEED)
        w, h = window
        x0, y0 = (W - w) // 2, (H - h) // 2
        prm = P.params(max_depth=depth, rng_mode=op.RNG_COUNTER_TABLE, seed=1, accum_mode=op.ACCUM_SUM, mesh_mode=op.MESH_NEAREST, threads=cores)

        def run(first, n):
            _, rays, secs = P.render_window(d, None, cam, table, W, H, x0, y0, w, h, first, n, prm)
            return rays, secs
        where = "centred %dx%d window of the %dx%d frame, brute force over all primitives" % (w, h, W, H)
    else:
        img = np.zeros((H, W, 4), np.float32)
        try:
            R, kind = op.RefOracle("fast"), "reference"
            table = R.make_table(TABLE_SEED)

            def run(first, n):
                _, rays, secs = R.render(d, tex, cam, table, W, H, first, n, depth=depth, threads=cores, image=img)
                return rays, secs
        except (FileNotFoundError, OSError):
            P, kind = op.PortOracle(), "port"
            table = P.make_table(TABLE_SEED)

            def run(first, n):
                _, rays, secs = P.render(d, tex, cam, table, W, H, first, n, P.params(max_depth=depth, threads=cores), image=img)
                return rays, secs
        where = "full %dx%d frame" % (W, H)
    k = 0
    for _ in range(warmup):
        run(k, spp_per_step); k += spp_per_step
    rays_t, secs_t = 0, 0.0
    for _ in range(steps):
        r, s = run(k, spp_per_step); k += spp_per_step
        rays_t += r; secs_t += s
    sample = ("%s, %d spp per step x %d steps (one pass per sample, like render()/renderAgain()), depth %d, %d host threads, %.1f s of CPU work"
              % (where, spp_per_step, steps, depth, cores, secs_t))
    return dict(rate=rays_t / secs_t / 1e6, cores=cores, kind=kind, rays=rays_t, secs=secs_t, sample=sample)


def reference_sampling(cfg):
    """(spp per step, window) of the bounded CPU sample of a config: sized for a few seconds of all host cores per step."""
    return {"c1": (16, None), "c2": (8, None), "c3": (8, None), "c4": (1, (32, 24)), "c5": (1, (128, 128))}[cfg["name"]]


def main_reference(a, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    spp_step, window = reference_sampling(cfg)
    r = cpu_reference_rate(cfg, spp_step, a.steps, a.warmup, window)
    rate = round(r["rate"], 3)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "Mrays/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(1e3 * r["secs"] / max(a.steps, 1), 3),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_block(cfg, spp_total=cfg["spp"], spp_per_gpu=None, arith="reference", rng="table", parallelism="host CPU cores",
                               rays_per_step=int(r["rays"] / max(a.steps, 1)), path_heads=False, bvh_nodes=0, queries=None,
                               l2="n/a (CPU arm)"),
        "cpu_baseline": {"value": rate, "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
        "e2e": {"value": rate, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), file=JSON_OUT, flush=True)


def config_block(cfg, spp_total, spp_per_gpu, arith, rng, parallelism, rays_per_step, path_heads, bvh_nodes, queries, l2):
    """Same keys in both arms.  The metric is a RATE: the reference arm times a bounded sample (reference_sample) of the same
    frame, scene, camera, depth and table seed; the B200 arm times the whole workload."""
    spp_step, window = reference_sampling(cfg)
    return {"workload": cfg["workload"], "spp_total": spp_total, "spp_per_gpu": spp_per_gpu, "arith": arith, "rng": rng,
            "parallelism": parallelism, "l2": l2, "rays_per_step": rays_per_step, "ray": RAY, "path_heads": path_heads,
            "queries_executed_per_step_rank0": queries, "bvh_nodes": bvh_nodes,
            "reference_sample": "the reference arm (bench.py --impl reference) renders %d spp per step of %s and reports a rate"
                                % (spp_step, "the full frame" if window is None else "a centred %dx%d window" % window)}


# ------------------------------------------------------------------------------------------------ B200 arm
def run_gpu_config(cfg, a, spp, steps, warmup, world, rank, local, with_clocks, with_cpu_baseline):
    """One BASELINE config on `world` GPUs -> the JSON object of that config (rank 0; None elsewhere)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    sys.path.insert(0, PKG)
    import rtb200
    from rtb200.multigpu import PeerFinalize, SharedHostImage, reduce_to_rank0, sample_range, split_image_check

    W, H, depth = cfg["W"], cfg["H"], cfg["depth"]
    arith = rtb200.ARITH_FAST if a.arith == "fast" else rtb200.ARITH_STRICT
    rng = {"counter": rtb200.RNG_COUNTER, "table": rtb200.RNG_TABLE, "counter_table": rtb200.RNG_COUNTER_TABLE}[a.rng]
    arr, tex, cam = load_scene(cfg)
    table = rtb200.make_random_table(TABLE_SEED)
    ctx = rtb200.Context(W, H, local)
    ctx.set_params(max_depth=depth, rng_mode=rng, seed=1, arith=arith, mesh_mode=cfg["mesh_mode"],
                   path_head=rtb200.HEAD_OFF if a.no_heads else rtb200.HEAD_ON)
    ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    fused = world > 1 and a.reduce != "nccl"
    accum = None
    if not fused:          # torch-owned buffer (what dist.reduce needs); the fused path exports the context's own buffer
        accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
        ctx.set_accum_device_buffer(accum.data_ptr(), accum.numel() * 4)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2
    peers = PeerFinalize(ctx, stream_ordered=True, device_barrier=(a.reduce == "fused")) if fused else None
    if world > 1 and a.split == "pixels":
        first, n_mine = 0, spp
        ctx.set_pixel_split(world, rank)
    else:
        first, n_mine = sample_range(spp, world, rank)

    def exchange(total_spp):
        if fused:
            peers.finalize(total_spp)
        else:
            reduce_to_rank0(accum)
            if rank == 0:
                ctx.finalize(total_spp)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: the picture of the real multi-process path (IPC peer buffers / NCCL) against rank 0 alone, before anything is timed
    check = split_image_check(ctx, cam, min(spp, 16 * world), world, rank, exchange, split=a.split) if world > 1 else None

    queries_steps = []

    def step():
        if fused:
            ctx.clear_accum()
        else:
            accum.zero_()
        ctx.invalidate_caches()          # every step rebuilds the path heads: nothing is carried over between steps
        ctx.render_samples(cam, first, n_mine, rtb200.ACCUM_SUM)
        st = ctx.stats()
        queries_steps.append(st.queries_last)
        exchange(spp)
        return st.rays_last, st.ms_trace_last

    for _ in range(warmup):
        step(); flush.zero_()
    launches0 = ctx.stats().kernel_launches
    sampler = ClockSampler(local) if (with_clocks and rank == 0) else None
    if sampler:
        sampler.start()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    rays_steps, kernel_ms = [], []
    for _ in range(steps):
        r, kms = step()
        rays_steps.append(r); kernel_ms.append(kms)
        flush.zero_()
    e1.record(stream)
    sync_all()
    elapsed_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    launches = ctx.stats().kernel_launches - launches0
    queries_timed = sum(queries_steps[-steps:])

    t = torch.tensor([elapsed_ms, float(sum(rays_steps)), float(launches), float(queries_timed)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        elapsed_ms, rays_total, launches, queries_total = tmax[0].item(), tsum[1].item(), int(tsum[2].item()), tsum[3].item()
    else:
        rays_total, queries_total = t[1].item(), t[3].item()
    value = rays_total / (elapsed_ms * 1e-3) / 1e6
    value_executed = queries_total / (elapsed_ms * 1e-3) / 1e6

    # ---- counted work of the LBVH kernel: the same launch through the instrumented copy (untimed)
    work = None
    bvh = int(ctx.stats().bvh_nodes) > 0
    if bvh and a.arith == "fast" and not a.no_heads:
        ctx.set_work_counting(True)
        ctx.clear_accum() if fused else accum.zero_()
        ctx.render_samples(cam, first, n_mine, rtb200.ACCUM_SUM)
        work = ctx.work_counts()
        ctx.set_work_counting(False)

    # ---- end to end through the C ABI with HOST buffers: every step uploads scene, textures, table (pinned host memory ->
    # device), renders, and the exchange / finalize step stores the picture straight into the caller's page-locked HOST image
    # (rt_set_finalize_target; N > 1: one shared host image, every rank writes its rows over its own PCIe link).
    shared = a.host_image == "shared" or world == 1
    host_image = SharedHostImage(H, W, world if shared else 1, rank)
    img_host = host_image.array
    tex_pinned = torch.from_numpy(tex).pin_memory()
    table_pinned = torch.from_numpy(table).pin_memory()
    h2d = tex_pinned.numel() * 4 + table_pinned.numel() * 4 + sum(v.nbytes for v in arr.values()) + 48
    d2h = img_host.nbytes                        # the whole picture crosses PCIe once per step

    e2e_parts = {"upload": 0.0, "render_exchange": 0.0, "readback": 0.0}

    def e2e_step():
        t_a = time.perf_counter()
        ctx.upload_textures(tex_pinned.numpy()); ctx.upload_scene(arr); ctx.set_random_table(table_pinned.numpy())
        t_b = time.perf_counter()
        r, _ = step()
        if shared:
            stream.synchronize()                 # rank 0: behind the last rank barrier, i.e. every rank's rows are in host memory
        t_c = time.perf_counter()
        if not shared and rank == 0:
            ctx.read_image(img_host)
        t_d = time.perf_counter()
        e2e_parts["upload"] += t_b - t_a; e2e_parts["render_exchange"] += t_c - t_b; e2e_parts["readback"] += t_d - t_c
        return r

    if shared:
        ctx.set_finalize_target(host_image.ptr, host_image.nbytes)
    e2e_step()
    sync_all()
    e2e_image_ok = bool(img_host[..., 3].min() == 1.0 and img_host[..., :3].max() > 0.1) if rank == 0 else True
    for k_ in e2e_parts:
        e2e_parts[k_] = 0.0
    t0 = time.perf_counter()
    e2e_rays = 0
    # enough steps that the fixed cost of the closing device + rank synchronisation (a few ms) does not pose as per-step time
    e2e_steps = max(steps, 10) if elapsed_ms / steps < 100.0 else max(1, min(steps, 3))
    for _ in range(e2e_steps):
        e2e_rays += e2e_step()
    sync_all()
    e2e_s = time.perf_counter() - t0
    ctx.set_finalize_target(None, 0)
    t2 = torch.tensor([e2e_s, float(e2e_rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        m = t2.clone(); dist.all_reduce(m, op=dist.ReduceOp.MAX)
        s = t2.clone(); dist.all_reduce(s, op=dist.ReduceOp.SUM)
        e2e_s, e2e_rays = m[0].item(), s[1].item()
    e2e_value = e2e_rays / e2e_s / 1e6

    # ---- the reference's own call pattern through the C ABI (main.cpp:91-95 -> src/raytracer.cpp:127-165): render() once, then
    # renderAgain() per sample, the image read at the end.  rt_render / rt_render_again record the samples and one launch renders
    # them when the image is read (bit-identical to one launch per call); RTB200_EAGER=1 is the reference's schedule, one launch
    # and one host round trip per sample, timed on fewer samples.
    api = None
    if world == 1 and cfg["name"] in ("c1", "c2"):
        def api_run(n_samples):
            ctx.invalidate_caches()
            t_a = time.perf_counter()
            ctx.render(cam)
            for _ in range(n_samples - 1):
                ctx.render_again(cam)
            ctx.read_image(img_host)
            dt = time.perf_counter() - t_a
            return ctx.stats().rays_total / dt / 1e6, dt
        api_run(min(spp, 64))
        v_def, s_def = api_run(spp)
        os.environ["RTB200_EAGER"] = "1"
        n_eager = min(spp, 128)
        api_run(8)
        v_eag, s_eag = api_run(n_eager)
        del os.environ["RTB200_EAGER"]
        api = {"value": round(v_def, 1), "unit": "Mrays/s", "samples": spp, "seconds": round(s_def, 4),
               "what": "rt_render + %d x rt_render_again + rt_read_image (host wall clock, image copied to the host): what RayTracer::render() / "
                       "renderAgain() of include/raytracer.h call; the recorded samples are rendered by one launch when the image is read" % (spp - 1),
               "eager": {"value": round(v_eag, 1), "samples": n_eager, "us_per_sample": round(1e6 * s_eag / n_eager, 1),
                         "what": "RTB200_EAGER=1: one launch + one host synchronisation per sample, the reference's schedule (src/raytracer.cpp:137-140,158-161)"}}

    out = None
    if rank == 0:
        k_ms = sum(kernel_ms) / len(kernel_ms)
        rays_launch, queries_launch = rays_steps[-1], queries_steps[steps + warmup - 1]
        kernel = ("k_trace_tail_%s_bvh" if bvh else ("k_trace_%s" if a.no_heads else "k_trace_tail_%s")) % a.arith
        if a.no_heads and bvh:
            kernel = "k_trace_%s_bvh" % a.arith
        peak_gslots = PEAK["gslots"]
        if work is not None:         # BVH rule of SURVEY.md 8(d): counted work of the launch
            slots_launch = (SLOT_BOX * 2.0 * work["node_visits"] + SLOT_SPHERE * work["sphere_tests"] + SLOT_LENS * work["lens_tests"]
                            + SLOT_TRI * work["triangle_tests"] + SLOT_PLANE * work["plane_tests"] + SLOT_RAY * work["queries"])
            slots_per_query = slots_launch / max(work["queries"], 1)
        elif cfg["slots"] is not None and not bvh:
            slots_per_query = cfg["slots"]
            slots_launch = slots_per_query * queries_launch
        else:
            slots_per_query = slots_launch = None
        executed = slots_launch / (k_ms * 1e-3) / 1e9 if slots_launch is not None else None
        hbm_peak = PEAKS_FILE.get("hbm_gbs", 6650.0)
        alg_bytes = 2.0 * W * H * 16          # accumulation buffer read + write, once per launch
        traffic = NCU_TRAFFIC.get((cfg["name"], str(n_mine))) if (a.arith == "fast" and not a.no_heads) else None
        roofline = {
            "bound": "fp32_issue", "achieved": round(executed, 1) if executed is not None else None, "peak": round(peak_gslots, 1),
            "unit": "G lane-slots/s",
            "frac": round(executed / peak_gslots, 4) if (executed is not None and peak_gslots) else None,
            "traffic": traffic,
            "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum of one launch of this kernel at this shape, bytes, from the ncu --set full "
                            "captures summarised under profiles/ (r02_ncu_traffic.json); null where no capture of this shape exists",
            "kernel": kernel, "kernel_ms": round(k_ms, 3),
            "algorithmic_gain": round(rays_launch / max(queries_launch, 1), 4),
            "note": "achieved = FP32 issue slots of the work the kernel EXECUTED in one launch / its duration (CUDA events on the launching "
                    "stream); frac = achieved / measured peak.  algorithmic_gain = reference path segments / executed queries: the head of a "
                    "pixel's path is traced once per pixel instead of once per sample, which is why value exceeds value_executed",
            "slots_per_query": round(slots_per_query, 1) if slots_per_query is not None else None,
            "slots_source": SLOTS_SOURCE[None] if work is not None else SLOTS_SOURCE.get(cfg["slots"]),
            "queries_per_launch": int(queries_launch), "segments_per_launch": int(rays_launch),
            "peak_source": "measured live: dependent-free FFMA loop on all SMs (rt_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 entry",
            "hbm": {"achieved_gbs": round(alg_bytes / (k_ms * 1e-3) / 1e9, 2), "peak_gbs": hbm_peak,
                    "measured_gbs": round(traffic / (k_ms * 1e-3) / 1e9, 2) if traffic else None,
                    "note": "evidence only: the accumulation buffer read + written once per launch is the algorithmic traffic, not the bound"},
        }
        if work is not None:
            q = max(work["queries"], 1)
            roofline["counted_work"] = dict(work, node_visits_per_query=round(work["node_visits"] / q, 2),
                                            leaf_tests_per_query=round((work["sphere_tests"] + work["lens_tests"] + work["triangle_tests"]) / q, 2))
        out = {
            "metric": METRIC, "value": round(value, 1), "value_executed": round(value_executed, 1), "unit": "Mrays/s", "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": round(elapsed_ms / steps, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_block(cfg, spp_total=spp, spp_per_gpu=n_mine, arith=a.arith, rng=a.rng,
                                   parallelism=("%s x%d + %s" % ("interleaved pixel-group split (all samples of a pixel on one GPU)" if a.split == "pixels" else "spp-split", world, ("fused peer-memory reduce+finalize, device-side rank barrier (rt_finalize_peers_barrier_async)" if a.reduce == "fused"
                                                                              else "fused peer-memory reduce+finalize between two NCCL barriers") if fused else "1 NCCL reduce"))
                                   if world > 1 else "single GPU",
                                   rays_per_step=int(rays_total / steps), path_heads=not a.no_heads, bvh_nodes=int(ctx.stats().bvh_nodes),
                                   queries=int(queries_launch),
                                   l2="flushed between steps (256 MiB memset); the scene is re-read from HBM by the first blocks of every step"),
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "host_image_complete": e2e_image_ok,
                    "rank0_ms_per_step": {k_: round(1e3 * v_ / e2e_steps, 3) for k_, v_ in e2e_parts.items()},
                    "what": "rt_upload_textures + rt_upload_scene + rt_set_random_table from pinned host memory, rt_render_samples, exchange + "
                            "finalize " + ("storing the picture straight into the caller's page-locked host image (rt_set_finalize_target)" if shared
                                         else "into rank 0's device image, rt_read_image into rank 0's pinned host buffer") + ", every step"},
            "gpu_launches": int(launches),
            "roofline": roofline,
        }
        if check is not None:
            out["multi_gpu_check"] = check
        if api is not None:
            out["api_path"] = api
        if world == 1 and with_cpu_baseline:
            spp_step, window = reference_sampling(cfg)
            reps = {"c1": 2, "c2": 4, "c3": 3, "c4": 1, "c5": 2}[cfg["name"]]
            r = cpu_reference_rate(cfg, spp_step * (4 if cfg["name"] == "c2" else 1), reps, 1 if window is None else 0, window)
            out["cpu_baseline"] = {"value": round(r["rate"], 3), "unit": "Mrays/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
    if peers is not None:
        if a.reduce == "fused" and ctx.peer_barrier_timed_out():
            raise SystemExit("bench.py: a device-side rank barrier gave up waiting (a rank never arrived): the numbers of this run are void")
        peers.close()
    ctx.close()
    host_image.close()
    del flush, accum
    torch.cuda.empty_cache()
    return out


def pin_to_gpu_numa_node(gpu_index):
    """One process per GPU: run this rank's host threads on the CPU cores next to its GPU (NVML's CPU affinity of the device), so
    that the page-locked staging buffers it allocates are first touched on that NUMA node and its PCIe copies do not cross the
    socket interconnect.  Best effort: without NVML, or where the mask is empty, nothing changes."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


PEAK = {"gslots": 0.0, "mhz": 0.0}
PEAKS_FILE = {}


def main_b200(a, cfg):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, PKG)
    import rtb200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    if world > 1:
        pin_to_gpu_numa_node(local)
    torch.cuda.set_device(local)
    # a real stream shared by the context (rt_set_stream), torch's fills and the NCCL collectives: torch's default stream is the
    # null handle, which rt_set_stream reads as "the context's private stream"
    torch.cuda.set_stream(torch.cuda.Stream())
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    PEAK["gslots"], PEAK["mhz"] = rtb200.measure_fp32_peak(local)
    try:
        PEAKS_FILE.update(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))))
    except Exception:
        pass
    spp = a.spp if a.spp is not None else cfg["spp"]
    out = run_gpu_config(cfg, a, spp, a.steps, a.warmup, world, rank, local, with_clocks=True, with_cpu_baseline=not a.no_cpu_baseline)
    extras = []
    if cfg["name"] == "c2" and not a.no_extras and a.spp is None and not a.no_heads:
        # the other BASELINE configs, briefly (3 warm-up + 2 timed steps each at a reduced sample count where the full one takes
        # seconds; the rate does not depend on it), so that the driver's BENCH / SCALE records carry them at every N
        for name in ("c1", "c3", "c4", "c5"):
            c = CONFIGS[name]
            e = run_gpu_config(c, a, max(c["short_spp"], world), 2, 3, world, rank, local, with_clocks=False,
                               with_cpu_baseline=not a.no_cpu_baseline)
            if e is not None:
                extras.append({k: e[k] for k in ("value", "value_executed", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "config", "e2e",
                                                 "gpu_launches", "roofline", "multi_gpu_check", "cpu_baseline", "api_path") if k in e})
    if rank == 0:
        if extras:
            out["extra_configs"] = extras
        print(json.dumps(out), file=JSON_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    JSON_OUT = claim_stdout()
    config = dict(CONFIGS[args.config], name=args.config)
    for _n, _c in CONFIGS.items():
        _c["name"] = _n
    if args.impl == "reference":
        main_reference(args, config)
    else:
        main_b200(args, config)
