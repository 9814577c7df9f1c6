"""tests/golden/make_fullsize_blockmeans.py — full-resolution golden statistics of BASELINE configs C2 and C3, produced
by the reference's own kernel text compiled as C++ (oracle/_ref/libref_fast.so).  Run in the BUILD container only:

    python tests/golden/make_fullsize_blockmeans.py

Writes golden_fullsize.npz: per config the 16x16 block means of the LINEAR radiance (image**2) at 1920x1080 after
SPP samples for two different random tables (seeds 12345 and 999), the ray counts, and the primary-id histogram.
The difference between the two seeds calibrates the Monte-Carlo noise the GPU tests may show against it; the mean
of the two is the golden estimate.  About 5 minutes of CPU time on 8 cores."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
import oracle_py as op  # noqa: E402
from rtb200 import fixtures as fx  # noqa: E402

W, H, SPP, B = 1920, 1080, 192, 16


def block_means(img, b=B):
    h, w = img.shape[:2]
    lin = img[: h // b * b, : w // b * b, :3].astype(np.float64) ** 2
    return lin.reshape(h // b, b, w // b, b, 3).mean((1, 3)).astype(np.float32)


def main():
    R = op.RefOracle("fast")
    out = dict(W=W, H=H, spp=SPP, block=B, build=R.build_info())
    for name, loader in (("c2", fx.default_scene), ("c3", fx.c3_scene)):
        arr, tex8 = loader()
        tex = fx.textures_to_float(tex8)
        d = op.make_desc(arr)
        cam = fx.camera(60, W / H)
        for seed in (12345, 999):
            img, rays, secs = R.render(d, tex, cam, R.make_table(seed), W, H, 0, SPP, depth=30)
            out["%s_blocks_seed%d" % (name, seed)] = block_means(img)
            out["%s_rays_seed%d" % (name, seed)] = rays
            print(name, seed, rays, "%.1f s" % secs, flush=True)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_fullsize.npz"), **out)


if __name__ == "__main__":
    main()
