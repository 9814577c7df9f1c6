"""GPU: host-side cost of rt_upload_scene and of the first render (LBVH build) on the C4 scene."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import numpy as np
import rtb200 as rt
from rtb200 import fixtures as fx, synth
arr, _ = synth.c4_scene(subdiv=8)
W, H = 3840, 2160
cam = fx.camera(60, W / H); table = rt.make_random_table(12345)
with rt.Context(W, H) as ctx:
    ctx.set_params(max_depth=30, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, mesh_mode=1)
    ctx.set_random_table(table)
    for i in range(3):
        t0 = time.perf_counter(); ctx.upload_scene(arr); t1 = time.perf_counter()
        ctx.render_samples(cam, 0, 1, rt.ACCUM_SUM); t2 = time.perf_counter()
        ctx.render_samples(cam, 1, 1, rt.ACCUM_SUM); t3 = time.perf_counter()
        print("upload_scene %.1f ms, first render (LBVH build + heads) %.1f ms, second render %.1f ms, bvh nodes %d" %
              (1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), ctx.stats().bvh_nodes), flush=True)
    ids, _ = ctx.primary_ids(cam)
    import hashlib
    print("primary ids sha", hashlib.sha256(ids.tobytes()).hexdigest()[:16])
