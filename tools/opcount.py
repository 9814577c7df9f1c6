"""FP32 issue slots per path segment of the reference algorithm, measured by running the reference's own
kernel text with an operation-counting float (oracle/ref_opcount.cpp -> oracle/_ref/libref_opcount.so).
usage: python tools/opcount.py [default|c3] [W H spp depth]     (CPU only, one thread; test / measurement infrastructure)"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import oracle_py as op
import scene_py as sp
from rtb200 import fixtures as fx

FLOAT_FIELDS = {"materials": (sp.MATERIAL_DT, ["color", "extra"]), "spheres": (sp.SPHERE_DT, ["pos", "r"]),
                "planes": (sp.PLANE_DT, ["pos", "normal"]), "lenses": (sp.LENS_DT, ["pos", "p1", "p2", "r1", "r2"])}


def untag(a):
    """clear the lowest mantissa bit (the op counter's 'unconsumed product' tag) of finite non-zero floats"""
    a = np.ascontiguousarray(a, dtype=np.float32).copy()
    u = a.view(np.uint32)
    ok = np.isfinite(a) & (a != 0)
    u[ok] &= np.uint32(0xFFFFFFFE)
    return a


def count(name="default", W=480, H=270, spp=4, depth=30, table_seed=12345):
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_opcount.so"))
    arr, tex8 = fx.c3_scene() if name == "c3" else fx.default_scene()
    arr = {k: np.ascontiguousarray(v).copy() for k, v in arr.items()}
    for k, (dt, fields) in FLOAT_FIELDS.items():
        rec = arr[k].reshape(-1).view(dt)
        for f in fields:
            rec[f] = untag(rec[f])
    arr["vertices"] = untag(arr["vertices"].view(np.float32)).view(arr["vertices"].dtype).reshape(arr["vertices"].shape)
    arr["uvs"] = untag(arr["uvs"].view(np.float32)).view(arr["uvs"].dtype).reshape(arr["uvs"].shape)
    tex = untag(fx.textures_to_float(tex8))
    cam = untag(fx.camera(60, W / H))
    table = untag(op.PortOracle().make_table(table_seed))
    d = op.make_desc(arr)
    img = np.zeros((H, W, 4), np.float32)
    counts = (C.c_uint64 * 9)()
    fp = lambda a: a.ctypes.data_as(C.c_void_p)
    L.ref_opcount_render.restype = None
    L.ref_opcount_render(C.byref(d), fp(tex), tex.shape[2], tex.shape[1], tex.shape[0], fp(cam), fp(table), fp(img),
                         W, H, 0, 0, W, H, 0, spp, depth, counts)
    add, mul, cmp_, div, sqrt, pow_, dmul, fus, rays = [int(c) for c in counts]
    total = add + mul + cmp_ + div + sqrt + pow_ + dmul
    per = lambda v: v / rays
    return {"scene": name, "W": W, "H": H, "spp": spp, "depth": depth, "rays": rays, "rays_per_pixel_sample": rays / (W * H * spp),
            "per_ray": {"add_sub": per(add), "mul": per(mul), "compare": per(cmp_), "div": per(div), "sqrt": per(sqrt),
                        "pow": per(pow_), "double_mul": per(dmul), "fma_fusable_pairs": per(fus)},
            "ops_per_ray": per(total), "slots_per_ray": per(total - fus)}


if __name__ == "__main__":
    import json
    a = sys.argv[1:]
    kw = {}
    if len(a) >= 5:
        kw = dict(W=int(a[1]), H=int(a[2]), spp=int(a[3]), depth=int(a[4]))
    print(json.dumps(count(a[0] if a else "default", **kw), indent=1))
