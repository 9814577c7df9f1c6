"""tests/golden/make_fixtures.py — regenerates the committed fixtures.  Run in the BUILD container only
(needs /root/reference and oracle/_ref):   python tests/golden/make_fixtures.py

The reference ships no tests, golden vectors or fixtures for this path, so the vectors here
are outputs of the reference's own kernel text compiled as C++ (oracle/_ref/libref_strict.so) on inputs
prepared by oracle/scene_py.py from the reference's assets:

  (rtb200/data/) scene_default.npz   assets/scenes/scene.scene parsed into the reference's host AoS arrays + RGBA8 textures
  (rtb200/data/) scene_c3.npz        tests/golden/scenes/c3_cube.scene (reference cube meshes + die textures), config C3
  golden_default.npz  reference outputs on scene_default: 160x120 image after 1 and 4 samples (depth 30),
                      primary ids/t at 160x120 and ids at 800x600, ray counts, table checksum; single-bounce
                      vectors (ray in -> ray out, throughput) for 256 recorded rays
  golden_c3.npz       the same for scene_c3 at 160x90
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as op  # noqa: E402
import scene_py as sp  # noqa: E402

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(os.path.dirname(os.path.dirname(HERE)), "opencl-raytracing_b200", "rtb200", "data")   # package data of rtb200
TABLE_SEED = 12345


def pack(arr):
    out = {}
    for k, v in arr.items():
        if k == "textures_rgba8":
            out[k] = v
        else:
            a = np.ascontiguousarray(v)
            out[k] = a.view(np.uint8).reshape(a.shape[0], -1) if a.size else np.zeros((0, max(a.dtype.itemsize if a.ndim == 1 else a.dtype.itemsize * a.shape[1], 1)), np.uint8)
    return out


def resolver(p):
    return os.path.join(REF, "assets", "textures", os.path.basename(p))


def bounce_vectors(R, d, tex, table, cam, W, H, n=256, seed=7):
    """Record rays along real paths (so every material type is visited) and their one-bounce results."""
    rng = np.random.RandomState(seed)
    rays_in, rays_out, cols, meta = [], [], [], []
    tries = 0
    while len(rays_in) < n and tries < 100000:
        tries += 1
        x, y = int(rng.randint(W)), int(rng.randint(H))
        llc, hor, ver = cam[3:6], cam[6:9], cam[9:12]
        v = llc + np.float32(np.float32(x) / np.float32(W)) * hor + np.float32(np.float32(y) / np.float32(H)) * ver
        v = (v / np.sqrt((v * v).sum())).astype(np.float32)
        ray = np.concatenate([cam[0:3], v]).astype(np.float32)
        sample = int(rng.randint(64))
        for b in range(6):
            r_out, col = R.bounce(d, tex, table, ray, b + sample, x, y)
            rays_in.append(ray.copy()); rays_out.append(r_out.copy()); cols.append(col.copy())
            meta.append((sample, b, x, y))
            if np.array_equal(r_out, ray) or len(rays_in) >= n:   # miss or light: ray unchanged
                break
            ray = r_out
    return (np.array(rays_in, np.float32), np.array(rays_out, np.float32), np.array(cols, np.float32),
            np.array(meta, np.uint32))


def golden(R, arr, tex, table, W, H, fov=60, pos=(0, 0, 0), yaw=0.0, pitch=0.0, big=None):
    d = op.make_desc(arr)
    cam = sp.camera_data(fov, W / H, pos, yaw, pitch)
    img1, rays1, _ = R.render(d, tex, cam, table, W, H, 0, 1, depth=30)
    img4, rays4, _ = R.render(d, tex, cam, table, W, H, 1, 3, depth=30, image=img1.copy())
    img8, rays8, _ = R.render(d, tex, cam, table, W, H, 0, 4, depth=8)
    ids, ts, bad = R.primary_ids(d, cam, W, H)
    assert bad == 0
    out = dict(W=W, H=H, cam=cam, image_1spp=img1[..., :3].copy(), image_4spp=img4[..., :3].copy(),
               image_4spp_depth8=img8[..., :3].copy(), rays_1spp=rays1, rays_4spp=rays1 + rays4,
               rays_4spp_depth8=rays8, ids=ids, ts=ts)
    if big:
        BW, BH = big
        camb = sp.camera_data(fov, BW / BH, pos, yaw, pitch)
        idsb, tsb, bad = R.primary_ids(d, camb, BW, BH)
        assert bad == 0
        out.update(big_W=BW, big_H=BH, big_cam=camb, big_ids=idsb)
    ri, ro, co, me = bounce_vectors(R, d, tex, table, cam, W, H)
    out.update(bounce_in=ri, bounce_out=ro, bounce_col=co, bounce_meta=me)
    return out


def main():
    R = op.RefOracle("strict")
    table = R.make_table(TABLE_SEED)
    tsum = hashlib.sha256(table.tobytes()).hexdigest()

    sc = sp.parse_scene_file(os.path.join(REF, "assets/scenes/scene.scene"), root=REF)
    sc.load_textures(resolver)
    arr = sc.arrays()
    np.savez_compressed(os.path.join(DATA, "scene_default.npz"), **pack(arr))
    tex = sp.texture_from_rgba8(arr["textures_rgba8"])
    g = golden(R, arr, tex, table, 160, 120, big=(800, 600))
    g.update(table_seed=TABLE_SEED, table_sha256=tsum)
    np.savez_compressed(os.path.join(HERE, "golden_default.npz"), **g)

    sc3 = sp.parse_scene_file(os.path.join(HERE, "scenes/c3_cube.scene"), root=REF)
    sc3.load_textures(resolver)
    arr3 = sc3.arrays()
    np.savez_compressed(os.path.join(DATA, "scene_c3.npz"), **pack(arr3))
    tex3 = sp.texture_from_rgba8(arr3["textures_rgba8"])
    g3 = golden(R, arr3, tex3, table, 160, 90, big=(480, 270))
    g3.update(table_seed=TABLE_SEED, table_sha256=tsum)
    np.savez_compressed(os.path.join(HERE, "golden_c3.npz"), **g3)
    for d_, f in ((DATA, "scene_default.npz"), (HERE, "golden_default.npz"), (DATA, "scene_c3.npz"), (HERE, "golden_c3.npz")):
        print(f, os.path.getsize(os.path.join(d_, f)))


if __name__ == "__main__":
    main()
