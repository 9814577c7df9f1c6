#!/usr/bin/env python
"""tools/sanitize_small.py — a few tiny renders for compute-sanitizer (memcheck / racecheck): brute-force fast arm with linear sums
(sample stealing), strict arm, and both arms through the LBVH (256-bit record loads).  usage: compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import numpy as np
import rtb200 as rt
from rtb200 import fixtures as fx, synth

table = rt.make_random_table(12345)
arr, tex8 = fx.default_scene()
tex = fx.textures_to_float(tex8)
W, H = 96, 64
cam = fx.camera(60, W / H)
with rt.Context(W, H) as ctx:
    ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
    for arith in (rt.ARITH_FAST, rt.ARITH_STRICT):
        ctx.set_params(max_depth=8, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, arith=arith)
        ctx.clear_accum(); ctx.render_samples(cam, 0, 24, rt.ACCUM_SUM)
        a = ctx.read_accum()
        print("brute", arith, float(a[..., :3].mean()), int(ctx.stats().rays_last), flush=True)
arrays, info = synth.scene(n_spheres=600, mesh_subdiv=3, seed=3)
with rt.Context(W, H) as ctx:
    ctx.upload_scene(arrays); ctx.set_random_table(table)
    for arith in (rt.ARITH_FAST, rt.ARITH_STRICT):
        ctx.set_params(max_depth=8, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, arith=arith, mesh_mode=rt.MESH_NEAREST, accel=rt.ACCEL_BVH)
        ctx.clear_accum(); ctx.render_samples(cam, 0, 8, rt.ACCUM_SUM)
        a = ctx.read_accum()
        print("lbvh", arith, float(a[..., :3].mean()), int(ctx.stats().rays_last), int(ctx.stats().bvh_nodes), flush=True)
print("done")
