"""BASELINE configs C4 (1 310 720 triangles + 1001 spheres) and C5 (100 001 spheres, depth 16) at 3840x2160 through the
GPU-built LBVH, beyond primary hits: a centred crop of the 4K frame (BASELINE.md §2's crop; rt_set_region = the same
work-items of the full NDRange, same primary rays and per-pixel RNG keys) rendered by the strict kernels equals the
oracle's O(N) brute force with the nearest-front-facing mesh rule bit for bit — image and ray count — with path heads on
and off; the fast arm's block means agree with the strict arm's within twice the seed-to-seed noise.

Crop sizes: the oracle's brute force costs 43 ms of one core per ray on C4 and 0.5 ms on C5 (measured), so C5 takes the
full 256 x 256 crop and C4 a centred 48 x 32 one (mesh silhouette + floor), 2 samples each."""
import numpy as np
import pytest

import oracle_py as op

pytestmark = pytest.mark.gpu
WB, HB = 3840, 2160


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def block_means(lin, b=16):
    h, w = lin.shape[:2]
    return lin[: h // b * b, : w // b * b, :3].astype(np.float64).reshape(h // b, b, w // b, b, 3).mean((1, 3))


@pytest.mark.parametrize("name,w,h,spp,depth", [("c5", 256, 256, 2, 16), ("c4", 48, 32, 2, 30)])
def test_crop_through_lbvh_equals_oracle_brute_force(port, table, name, w, h, spp, depth):
    import rtb200 as rt
    from rtb200 import fixtures as fx, synth
    arrays, info = synth.c4_scene(subdiv=8) if name == "c4" else synth.c5_scene()
    cam = fx.camera(60, WB / HB)
    d = op.make_desc(arrays)
    x0, y0 = (WB - w) // 2, (HB - h) // 2
    prm = port.params(max_depth=depth, rng_mode=op.RNG_COUNTER_TABLE, seed=1, accum_mode=op.ACCUM_SUM, mesh_mode=op.MESH_NEAREST)
    want, rays_want, secs = port.render_window(d, None, cam, table, WB, HB, x0, y0, w, h, 0, spp, prm)
    assert rays_want > 2 * w * h * spp                       # bounces happen inside the crop
    with rt.Context(WB, HB) as ctx:
        ctx.upload_scene(arrays)
        ctx.set_random_table(table)
        ctx.set_region(x0, y0, w, h)
        for head in (rt.HEAD_ON, rt.HEAD_OFF):
            ctx.set_params(max_depth=depth, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, arith=rt.ARITH_STRICT, mesh_mode=rt.MESH_NEAREST, path_head=head)
            ctx.clear_accum()
            ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
            st = ctx.stats()
            assert st.bvh_nodes == info["spheres"] + info["triangles"] - 1
            acc = ctx.read_accum()
            got = acc[y0:y0 + h, x0:x0 + w]
            n_diff = int((bits(got) != bits(want)).any(-1).sum())
            print("%s crop %dx%d, %d spp, heads %s: %d rays (oracle %d, %.1f s of brute force), %d differing pixels"
                  % (name, w, h, spp, "on" if head == rt.HEAD_ON else "off", st.rays_last, rays_want, secs, n_diff))
            assert n_diff <= 1                               # a powf double-rounding flip (p ~ 1e-8 per dielectric bounce)
            assert abs(int(st.rays_last) - int(rays_want)) <= 40 * n_diff
            outside = acc.copy()
            outside[y0:y0 + h, x0:x0 + w] = 0.0
            assert not outside.any()                         # nothing outside the region is touched
        # the reference's sqrt(mean) chain through the region
        ctx.set_params(max_depth=depth, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, arith=rt.ARITH_STRICT, mesh_mode=rt.MESH_NEAREST)
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        chain = ctx.read_image()[y0:y0 + h, x0:x0 + w, :3]
        assert np.abs(chain - np.sqrt(want[..., :3] / spp)).max() <= 2e-6


@pytest.mark.parametrize("name,depth", [("c4", 30), ("c5", 16)])
def test_crop_fast_arm_statistics(table, name, depth):
    """Fast arithmetic vs strict arithmetic (which the test above pins to the oracle) on the centred 256 x 256 crop: different
    rounding, hence different paths, same distribution — block means within twice the strict arm's seed-to-seed noise."""
    import rtb200 as rt
    from rtb200 import fixtures as fx, synth
    arrays, _ = synth.c4_scene(subdiv=8) if name == "c4" else synth.c5_scene()
    cam = fx.camera(60, WB / HB)
    w = h = 256
    x0, y0, spp = (WB - w) // 2, (HB - h) // 2, 64
    res = {}
    with rt.Context(WB, HB) as ctx:
        ctx.upload_scene(arrays)
        ctx.set_random_table(table)
        ctx.set_region(x0, y0, w, h)
        for key, arith, seed in (("strict_a", rt.ARITH_STRICT, 1), ("strict_b", rt.ARITH_STRICT, 2), ("fast", rt.ARITH_FAST, 3)):
            ctx.set_params(max_depth=depth, rng_mode=rt.RNG_COUNTER_TABLE, seed=seed, arith=arith, mesh_mode=rt.MESH_NEAREST)
            ctx.clear_accum()
            ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
            acc = ctx.read_accum()[y0:y0 + h, x0:x0 + w]
            assert np.all(acc[..., 3] == spp)
            res[key] = (block_means(acc[..., :3] / spp), ctx.stats().rays_last)
    rms = lambda a, b: float(np.sqrt(((a - b) ** 2).mean()))
    noise = rms(res["strict_a"][0], res["strict_b"][0])
    err = rms(res["fast"][0], res["strict_a"][0])
    print("%s crop: fast vs strict block-mean RMS %.5f, strict seed-to-seed %.5f" % (name, err, noise))
    assert err <= 2.0 * noise
    assert abs(res["fast"][0].mean() - res["strict_a"][0].mean()) <= 3.0 * noise
    assert abs(res["fast"][1] - res["strict_a"][1]) <= 0.02 * res["strict_a"][1]
