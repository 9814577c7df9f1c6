import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import numpy as np, oracle_py as op, rtb200
from rtb200 import fixtures as fx
P = op.PortOracle()
table = rtb200.make_random_table(12345)
for name, loader in (("default", fx.default_scene), ("c3", fx.c3_scene)):
    arr, tex8 = loader(); tex = fx.textures_to_float(tex8); d = op.make_desc(arr)
    for (W, H) in ((160, 120), (800, 600), (1920, 1080)):
        cam = fx.camera(60, W / H)
        want, wt = P.primary_ids(d, cam, W, H)
        with rtb200.Context(W, H) as ctx:
            ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
            for arith in (rtb200.ARITH_STRICT, rtb200.ARITH_FAST):
                ctx.set_params(arith=arith)
                ids, ts = ctx.primary_ids(cam)
                bad = np.argwhere(ids != want)
                print(name, W, H, "arith", arith, "mismatches", len(bad), "max|dt|", float(np.abs(ts - wt)[ids == want].max()))
                for (y, x) in bad[:8]:
                    print("   px", x, y, "want", hex(want[y, x]), wt[y, x], "got", hex(ids[y, x]), ts[y, x])
