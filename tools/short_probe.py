#!/usr/bin/env python
"""tools/short_probe.py — C2 rendered in short calls (what one rank of an N-way spp split does): ms per call for the block-order
variants.  usage: short_probe.py SPP"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import rtb200
from rtb200 import fixtures as fx
W, H = 1920, 1080
arr, tex8 = fx.default_scene()
tex = fx.textures_to_float(tex8)
cam = fx.camera(60, W / H)
table = rtb200.make_random_table(12345)
for spp in [int(x) for x in sys.argv[1].split(",")]:
    for env in ({}, {"RTB200_TAIL_LPT": "0"}, {"RTB200_TAIL_PILOT": "1"}, {"RTB200_TAIL_PILOT": "4"}, {"RTB200_TAIL_LPT": "0", "RTB200_TAIL_CHUNKS": "2"}):
        for k in ("RTB200_TAIL_CHUNKS", "RTB200_TAIL_PILOT", "RTB200_TAIL_LPT"):
            os.environ.pop(k, None)
        os.environ.update(env)
        with rtb200.Context(W, H, 0) as ctx:
            ctx.set_params(max_depth=30, rng_mode=rtb200.RNG_COUNTER_TABLE, seed=1, arith=rtb200.ARITH_FAST)
            ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
            best = None
            for rep in range(5):
                ctx.clear_accum(); ctx.invalidate_caches()
                ctx.render_samples(cam, 0, spp, rtb200.ACCUM_SUM)
                st = ctx.stats()
                if best is None or st.ms_last < best[0]:
                    best = (st.ms_last, st.ms_trace_last, st.queries_last, st.rays_last)
        print(json.dumps(dict(spp=spp, env=env, ms=round(best[0], 3), ms_trace=round(best[1], 3), grays_equiv=round(best[3] / best[0] / 1e6, 2))), flush=True)
