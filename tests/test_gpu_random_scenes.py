"""GPU parity on scenes the fixtures do not contain: the six seeded random scenes that pin the C oracle to the reference's
own kernel build on the CPU (tests/test_oracle_pinning.py::test_port_vs_reference_on_random_scenes — all six materials of
kernels/raytracer.cl:454-479, lenses :196-255, transformed meshes, random camera and depth) go through the strict CUDA
kernels: image bits, ray count, primary ids and t — with path heads on and off, in the table and the counter-table RNG mode,
through the reference's own call sequence and through the batched call.  The fast arm must keep the primary ids."""
import os
import shutil

import numpy as np
import pytest

import oracle_py as op

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def random_scene(seed, tmp_path):
    """Same generator and seeds as the CPU pinning test -> (arrays, float textures, camera, depth)."""
    import scene_py as sp
    from rtb200 import fixtures as fx
    from test_host_classes import ASSETS, _random_scene_text
    rng = np.random.default_rng(4000 + seed)
    for f in ("shapes.obj", "shapes.mtl", "tex_a.png", "tex_b.png"):
        shutil.copy(os.path.join(ASSETS, f), tmp_path / f)
    (tmp_path / "r.scene").write_text(_random_scene_text(rng))
    sc = sp.parse_scene_file(str(tmp_path / "r.scene"), root=str(tmp_path))
    sc.load_textures(lambda p: str(tmp_path / os.path.basename(p)))
    arr = sc.arrays()
    tex = sp.texture_from_rgba8(arr["textures_rgba8"])
    W, H = 96, 64
    depth = int(rng.integers(2, 12))
    cam = fx.camera(int(rng.integers(30, 90)), W / H, tuple(rng.uniform(-1, 1, 3)), float(rng.uniform(0, 360)), float(rng.uniform(-30, 30)))
    return arr, tex, cam, depth, W, H


@pytest.mark.parametrize("seed", range(6))
def test_random_scenes_strict_bit_exact(port, table, tmp_path, seed):
    import rtb200 as rt
    arr, tex, cam, depth, W, H = random_scene(seed, tmp_path)
    d = op.make_desc(arr)
    spp = 4
    want_ids, want_t = port.primary_ids(d, cam, W, H)
    observed = {}
    with rt.Context(W, H) as ctx:
        ctx.upload_textures(tex)
        ctx.upload_scene(arr)
        ctx.set_random_table(table)
        for rng_name, rng_gpu, rng_orc in (("table", rt.RNG_TABLE, op.RNG_TABLE), ("counter_table", rt.RNG_COUNTER_TABLE, op.RNG_COUNTER_TABLE)):
            want, rays_want, _ = port.render(d, tex, cam, table, W, H, 0, spp, port.params(max_depth=depth, rng_mode=rng_orc, seed=11))
            assert rays_want >= W * H * spp
            for head in (rt.HEAD_ON, rt.HEAD_OFF):
                ctx.set_params(max_depth=depth, rng_mode=rng_gpu, seed=11, arith=rt.ARITH_STRICT, path_head=head)
                ctx.render(cam)                                  # the reference's call sequence (src/raytracer.cpp:127-165)
                for _ in range(spp - 1):
                    ctx.render_again(cam)
                got = ctx.read_image()
                n_diff = int((bits(got) != bits(want)).any(-1).sum())
                observed[(rng_name, head)] = n_diff
                assert n_diff <= 1, (rng_name, head, n_diff)     # a powf double-rounding flip (p ~ 1e-8 per dielectric bounce)
                assert abs(int(ctx.stats().rays_total) - int(rays_want)) <= 20 * n_diff
                ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)   # the same samples in one launch
                assert np.array_equal(bits(ctx.read_image()), bits(got))
                assert abs(int(ctx.stats().rays_last) - int(rays_want)) <= 20 * n_diff
        ctx.set_params(max_depth=depth, arith=rt.ARITH_STRICT)
        ids, ts = ctx.primary_ids(cam)
        assert np.array_equal(ids, want_ids) and np.array_equal(bits(ts), bits(want_t))
        ctx.set_params(max_depth=depth, arith=rt.ARITH_FAST)
        ids_fast, _ = ctx.primary_ids(cam)
        assert np.array_equal(ids_fast, want_ids)
    print("random scene %d: differing pixels per (rng, head) = %s" % (seed, observed))


@pytest.mark.parametrize("seed", [0, 3])
def test_random_scenes_through_the_lbvh(port, table, tmp_path, seed):
    """The same scenes forced through the GPU-built LBVH (RT_ACCEL_BVH, nearest-front-facing mesh rule): strict image bits
    and ray count equal the oracle's brute force with the same mesh rule."""
    import rtb200 as rt
    arr, tex, cam, depth, W, H = random_scene(seed, tmp_path)
    d = op.make_desc(arr)
    spp = 3
    want, rays_want, _ = port.render(d, tex, cam, table, W, H, 0, spp,
                                     port.params(max_depth=depth, rng_mode=op.RNG_COUNTER_TABLE, seed=5, mesh_mode=op.MESH_NEAREST))
    with rt.Context(W, H) as ctx:
        ctx.upload_textures(tex)
        ctx.upload_scene(arr)
        ctx.set_random_table(table)
        for head in (rt.HEAD_ON, rt.HEAD_OFF):
            ctx.set_params(max_depth=depth, rng_mode=rt.RNG_COUNTER_TABLE, seed=5, arith=rt.ARITH_STRICT, mesh_mode=rt.MESH_NEAREST,
                           accel=rt.ACCEL_BVH, path_head=head)
            ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
            assert ctx.stats().bvh_nodes > 0
            n_diff = int((bits(ctx.read_image()) != bits(want)).any(-1).sum())
            assert n_diff <= 1, (head, n_diff)
            assert abs(int(ctx.stats().rays_last) - int(rays_want)) <= 20 * n_diff
