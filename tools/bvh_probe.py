#!/usr/bin/env python
"""tools/bvh_probe.py — quick timing + counted work of the LBVH configs (C4, C5) for A/B builds (RTB200_LIB=...).
Prints one JSON line per config: executed queries/s, reference-equivalent rays/s, node visits and leaf tests per query."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import numpy as np
import rtb200
from rtb200 import fixtures as fx, synth


def main():
    configs = sys.argv[1].split(",") if len(sys.argv) > 1 else ["c4", "c5"]
    spp = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    W, H = 3840, 2160
    cam = fx.camera(60, W / H)
    table = rtb200.make_random_table(12345)
    for name in configs:
        arr, _ = synth.c4_scene(subdiv=8) if name == "c4" else synth.c5_scene()
        depth = 30 if name == "c4" else 16
        with rtb200.Context(W, H, 0) as ctx:
            ctx.set_params(max_depth=depth, rng_mode=rtb200.RNG_COUNTER_TABLE, seed=1, arith=rtb200.ARITH_FAST, mesh_mode=rtb200.MESH_NEAREST)
            ctx.upload_scene(arr); ctx.set_random_table(table)
            ctx.clear_accum(); ctx.render_samples(cam, 0, 8, rtb200.ACCUM_SUM)
            best = None
            for rep in range(2):
                ctx.clear_accum(); ctx.invalidate_caches()
                ctx.render_samples(cam, 0, spp, rtb200.ACCUM_SUM)
                st = ctx.stats()
                r = dict(ms=st.ms_last, ms_trace=st.ms_trace_last, rays=st.rays_last, queries=st.queries_last)
                if best is None or r["ms"] < best["ms"]:
                    best = r
            ctx.set_work_counting(True)
            ctx.clear_accum()
            ctx.render_samples(cam, 0, 8, rtb200.ACCUM_SUM)
            wc = ctx.work_counts()
            ctx.set_work_counting(False)
            q = max(wc["queries"], 1)
            out = dict(config=name, lib=os.path.basename(rtb200.LIB_PATH), spp=spp, env={k: v for k, v in os.environ.items() if k.startswith("RTB200_") and k != "RTB200_LIB"},
                       grays_equiv=round(best["rays"] / best["ms"] / 1e6, 3), gqueries=round(best["queries"] / best["ms"] / 1e6, 3),
                       ms=round(best["ms"], 2), ms_trace=round(best["ms_trace"], 2),
                       nodes_per_query=round(wc["node_visits"] / q, 2), tri_per_query=round(wc["triangle_tests"] / q, 2),
                       sph_per_query=round(wc["sphere_tests"] / q, 2), flared_frac=round(wc["flared_queries"] / q, 4),
                       nodes_per_flared_query=round(wc["flared_node_visits"] / max(wc["flared_queries"], 1), 1), counts=wc)
            print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
