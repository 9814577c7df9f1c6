"""GPU tests at BASELINE.json's FULL sizes (C2: default scene 1920x1080 1024 spp; C3: cube scene 1920x1080), through
the C ABI.  The oracle cannot render these sizes in seconds, so they are checked
  * against committed statistics of the reference's OWN kernel at full resolution (tests/golden/golden_fullsize.npz,
    generator tests/golden/make_fullsize_blockmeans.py: 16x16 block means of the linear radiance after 192 samples for
    two random tables; their difference calibrates the Monte-Carlo noise), and
  * through size-independent properties: sample count in every pixel, split + sum == whole, strict arm == fast arm
    statistics, path heads on == off in segment count, segments per pixel-sample equal to the reference's.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_fullsize.npz")
W, H = 1920, 1080


def block_means(img, b=16):
    h, w = img.shape[:2]
    lin = img[: h // b * b, : w // b * b, :3].astype(np.float64) ** 2
    return lin.reshape(h // b, b, w // b, b, 3).mean((1, 3))


def rms(a, b):
    return float(np.sqrt(((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean()))


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def ctx_for(rt, scenes, name, table, **params):
    arr, tex = scenes[name]
    ctx = rt.Context(W, H)
    ctx.upload_textures(tex)
    ctx.upload_scene(arr)
    ctx.set_random_table(table)
    ctx.set_params(**params)
    return ctx


@pytest.mark.parametrize("name,scene,spp", [("c2", "default", 1024), ("c3", "c3", 1024)])
def test_full_size_against_reference_statistics(gold, scenes, table, name, scene, spp):
    """bench.py's configuration (fast arithmetic, counter-table RNG, path heads) at full size vs the reference kernel."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    cam = fx.camera(60, W / H)
    a, b = gold[name + "_blocks_seed12345"], gold[name + "_blocks_seed999"]
    noise = rms(a, b)                                   # two tables of the reference against each other, 192 spp
    rays_ref = float(gold[name + "_rays_seed12345"]) / (W * H * int(gold["spp"]))
    with ctx_for(rt, scenes, scene, table, max_depth=30, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=1) as ctx:
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        st = ctx.stats()
        ctx.finalize(spp)
        img = ctx.read_image()
        acc = ctx.read_accum()
    assert np.all(acc[..., 3] == spp) and np.all(img[..., 3] == 1.0)
    assert np.isfinite(img).all() and img.min() >= 0.0
    got = block_means(img)
    assert rms(got, a) <= 1.5 * noise, (rms(got, a), noise)
    assert abs(got.mean() - 0.5 * (a.mean() + b.mean())) <= 1.5e-3
    assert np.abs(got - a).max() <= 2.0 * np.abs(a.astype(np.float64) - b).max()
    # path segments per pixel-sample equal the reference's count (2.344 on C2) although only a part is executed
    assert abs(st.rays_last / (W * H * spp) - rays_ref) <= 2e-3 * rays_ref
    assert st.queries_last < 0.7 * st.rays_last


def test_full_size_properties_c2(scenes, table):
    """C2 at full size: strict and fast arms agree statistically, a two-way sample split sums to the whole, heads on/off
    trace the same number of segments (strict), and the reference accumulation chain equals sum -> finalize."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    cam = fx.camera(60, W / H)
    spp = 256
    kw = dict(max_depth=30, rng_mode=rt.RNG_COUNTER_TABLE, seed=4)
    with ctx_for(rt, scenes, "default", table, arith=rt.ARITH_FAST, **kw) as ctx:
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        whole = ctx.read_accum().copy()
        rays_fast = ctx.stats().rays_last
        ctx.finalize(spp)
        img_fast = ctx.read_image().copy()
        halves = np.zeros_like(whole)
        for first, n in ((0, 100), (100, 156)):
            ctx.clear_accum()
            ctx.render_samples(cam, first, n, rt.ACCUM_SUM)
            halves += ctx.read_accum()
        assert np.all(halves[..., 3] == spp)
        assert np.abs(halves - whole).max() <= 1e-4 * max(1.0, float(whole[..., :3].max()))
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)          # the reference's sqrt(mean) chain
        chain = ctx.read_image()
        assert np.abs(chain[..., :3] - img_fast[..., :3]).max() <= 2e-4
    with ctx_for(rt, scenes, "default", table, arith=rt.ARITH_STRICT, **kw) as ctx:
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        rays_on, q_on = ctx.stats().rays_last, ctx.stats().queries_last
        ctx.finalize(spp)
        img_strict = ctx.read_image().copy()
        ctx.set_params(arith=rt.ARITH_STRICT, path_head=rt.HEAD_OFF, **kw)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, 8, rt.ACCUM_SUM)
        rays8_off = ctx.stats().rays_last
        ctx.set_params(arith=rt.ARITH_STRICT, path_head=rt.HEAD_ON, **kw)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, 8, rt.ACCUM_SUM)
        assert ctx.stats().rays_last == rays8_off
    assert q_on < rays_on
    assert abs(rays_fast - rays_on) <= 2e-4 * rays_on
    # different arithmetic -> different paths, same distribution: the difference is Monte-Carlo noise of 256 spp
    assert rms(block_means(img_fast), block_means(img_strict)) <= 0.004
    assert abs(float((img_fast[..., :3] ** 2).mean()) - float((img_strict[..., :3] ** 2).mean())) <= 1e-3


@pytest.mark.parametrize("name,n_rays", [("c4", 800), ("c5", 4000)])
def test_full_size_lbvh_primary_hits_equal_oracle_brute_force(port, name, n_rays):
    """BASELINE configs C4 (1.31 M triangles + 1001 spheres) and C5 (100 001 spheres) at 3840x2160 through the GPU-built
    LBVH: the primary hit (id and t bits) of randomly chosen pixels equals the oracle's O(N) brute force — in the strict
    arm and, by construction (path heads are walked in Exact arithmetic), in the fast arm."""
    import oracle_py as op
    import rtb200 as rt
    from rtb200 import fixtures as fx, synth
    Wb, Hb = 3840, 2160
    arrays, info = synth.c4_scene(subdiv=8) if name == "c4" else synth.c5_scene()
    cam = fx.camera(60, Wb / Hb)
    d = op.make_desc(arrays)
    rng = np.random.RandomState(17)
    xs, ys = rng.randint(0, Wb, n_rays), rng.randint(Hb // 3, Hb, n_rays)      # lower two thirds: where the geometry is
    with rt.Context(Wb, Hb) as ctx:
        ctx.set_params(max_depth=16, rng_mode=rt.RNG_COUNTER, seed=1, arith=rt.ARITH_STRICT, mesh_mode=rt.MESH_NEAREST)
        ctx.upload_scene(arrays)
        ids, ts = ctx.primary_ids(cam)
        assert ctx.stats().bvh_nodes == info["spheres"] + info["triangles"] - 1
        ctx.set_params(max_depth=16, rng_mode=rt.RNG_COUNTER, seed=1, arith=rt.ARITH_FAST, mesh_mode=rt.MESH_NEAREST)
        ids_fast, _ = ctx.primary_ids(cam)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, 2, rt.ACCUM_SUM)
        acc = ctx.read_accum()
        st = ctx.stats()
    assert np.array_equal(ids_fast, ids)
    assert np.all(acc[..., 3] == 2) and np.isfinite(acc).all() and st.queries_last < st.rays_last
    llc, hor, ver = cam[3:6], cam[6:9], cam[9:12]
    kinds = set()
    for x, y in zip(xs, ys):
        v = llc + np.float32(np.float32(x) / np.float32(Wb)) * hor + np.float32(np.float32(y) / np.float32(Hb)) * ver
        # the kernels' primary direction is normalize() in the oracle's operation order; take the oracle's own
        hit, h = port.intersect(d, cam[0:3], _normalize_like_oracle(v), op.MESH_NEAREST)
        assert ids[y, x] == h.id, (x, y, hex(ids[y, x]), hex(h.id))
        if hit:
            assert np.float32(ts[y, x]).view(np.uint32) == np.float32(h.t).view(np.uint32)
            kinds.add(h.id >> 28)
    assert len(kinds) >= 2


def _normalize_like_oracle(v):
    """cl_shim / rt_oracle: normalize(v) = v * (1 / sqrt(dot(v, v))), dot = (x*x + y*y) + z*z, all binary32."""
    f = np.float32
    s = f(f(f(v[0] * v[0]) + f(v[1] * v[1])) + f(v[2] * v[2]))
    inv = f(1.0) / np.sqrt(s, dtype=np.float32)
    return np.array([v[0] * inv, v[1] * inv, v[2] * inv], np.float32)
