#!/usr/bin/env python
"""tools/split_probe.py — one GPU playing rank 0 of an N-way split of C2: kernel time and executed queries/s of the spp split
(whole frame, spp / N samples) against the pixel-group split (1 / N of the pixel groups, all samples)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import rtb200
from rtb200 import fixtures as fx

W, H, SPP = 1920, 1080, 1024
arr, tex8 = fx.default_scene()
tex = fx.textures_to_float(tex8)
cam = fx.camera(60, W / H)
table = rtb200.make_random_table(12345)


def run(tag, n_parts, part, first, n, env=None):
    for k in ("RTB200_TAIL_CHUNKS", "RTB200_HEAD_SORT", "RTB200_TAIL_LPT"):
        os.environ.pop(k, None)
    os.environ.update(env or {})
    with rtb200.Context(W, H, 0) as ctx:
        ctx.set_params(max_depth=30, rng_mode=rtb200.RNG_COUNTER_TABLE, seed=1, arith=rtb200.ARITH_FAST)
        ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
        ctx.set_pixel_split(n_parts, part)
        best = None
        for rep in range(4):
            ctx.clear_accum(); ctx.invalidate_caches()
            ctx.render_samples(cam, first, n, rtb200.ACCUM_SUM)
            st = ctx.stats()
            if best is None or st.ms_last < best[0]:
                best = (st.ms_last, st.ms_trace_last, st.queries_last, st.rays_last)
    print(json.dumps(dict(tag=tag, env=env or {}, ms=round(best[0], 3), ms_trace=round(best[1], 3), gq_per_s=round(best[2] / best[0] / 1e6, 2),
                          grays_equiv=round(best[3] / best[0] / 1e6, 2))), flush=True)


for N in (2, 8):
    run("spp split x%d: whole frame, %d spp" % (N, SPP // N), 1, 0, 0, SPP // N)
    run("pixel split x%d: 1/%d of the groups, %d spp" % (N, N, SPP), N, 0, 0, SPP)
    run("pixel split x%d, 1 chunk" % N, N, 0, 0, SPP, {"RTB200_TAIL_CHUNKS": "1"})
    run("pixel split x%d, chunks = N" % N, N, 0, 0, SPP, {"RTB200_TAIL_CHUNKS": str(N)})
    run("pixel split x%d, heads in slot order" % N, N, 0, 0, SPP, {"RTB200_HEAD_SORT": "0"})
    run("pixel split x%d, no block order" % N, N, 0, 0, SPP, {"RTB200_TAIL_LPT": "0"})
run("whole frame, 1024 spp", 1, 0, 0, SPP)
