"""Edge cases of the hot path on the GPU, each checked bit for bit against the oracle (strict arithmetic): empty and
ragged inputs, frame sizes that do not fill the 8x4 pixel tiles, depth 1, meshes longer than one 32-triangle chunk in
the brute-force path under both mesh rules, several mesh instances, zero-sample calls."""
import numpy as np
import pytest

import oracle_py as op

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def empty(n, item):
    return np.zeros((n, item), np.uint8)


def base_arrays(materials, spheres=None, planes=None):
    from rtb200 import synth
    return {"materials": synth._bytes(materials), "spheres": synth._bytes(spheres) if spheres is not None else empty(0, 32),
            "planes": synth._bytes(planes) if planes is not None else empty(0, 48), "lenses": empty(0, 64),
            "vertices": empty(0, 16), "uvs": empty(0, 8), "indices": empty(0, 4), "meshes": empty(0, 16), "models": empty(0, 12)}


def render_both(port, arrays, W, H, spp, depth, mesh_mode=0, accel=0, rng=1, table=None, cam=None, tex=None):
    import rtb200 as rt
    from rtb200 import fixtures as fx
    cam = fx.camera(60, W / H) if cam is None else cam
    prm = port.params(max_depth=depth, rng_mode=rng, seed=13, mesh_mode=mesh_mode)
    want, rays_want, _ = port.render(op.make_desc(arrays), tex, cam, table, W, H, 0, spp, prm)
    with rt.Context(W, H) as ctx:
        ctx.set_params(max_depth=depth, rng_mode=rng, seed=13, arith=rt.ARITH_STRICT, mesh_mode=mesh_mode, accel=accel)
        if tex is not None:
            ctx.upload_textures(tex)
        if table is not None:
            ctx.set_random_table(table)
        ctx.upload_scene(arrays)
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        got = ctx.read_image()
        st = ctx.stats()
    return got, want, st, rays_want


def test_scene_without_primitives_is_black(port):
    from rtb200 import synth
    arrays = base_arrays(synth.materials([(synth.T_DIFFUSE, (1, 1, 1), 1.0)]))
    got, want, st, rays_want = render_both(port, arrays, 37, 21, 3, 30)
    assert np.array_equal(bits(got), bits(want)) and not got[..., :3].any() and np.all(got[..., 3] == 1.0)
    assert st.rays_last == rays_want == 37 * 21 * 3          # one segment (the miss) per pixel-sample
    assert st.queries_last == 37 * 21                         # traced once per pixel: the whole frame is path heads


@pytest.mark.parametrize("W,H", [(1, 1), (7, 3), (9, 5), (64, 1), (1, 33), (250, 131)])
def test_ragged_frame_sizes(port, scenes, table, W, H):
    arr, tex = scenes["default"]
    got, want, st, rays_want = render_both(port, arr, W, H, 4, 30, rng=2, table=table, tex=tex)
    assert np.array_equal(bits(got), bits(want))
    assert st.rays_last == rays_want


@pytest.mark.parametrize("depth", [1, 2])
def test_minimum_depth(port, scenes, table, depth):
    """depth 1: every path is its first segment; the loop ends with the throughput after one material (cl:447-485)."""
    arr, tex = scenes["c3"]
    got, want, st, rays_want = render_both(port, arr, 160, 90, 3, depth, rng=0, table=table, tex=tex)
    assert np.array_equal(bits(got), bits(want)) and st.rays_last == rays_want
    if depth == 1:
        assert st.rays_last == 160 * 90 * 3


def test_zero_samples_is_a_no_op(scenes, table):
    import rtb200 as rt
    from rtb200 import fixtures as fx
    arr, tex = scenes["default"]
    cam = fx.camera(60, 2.0)
    with rt.Context(64, 32) as ctx:
        ctx.upload_textures(tex); ctx.upload_scene(arr); ctx.set_random_table(table)
        ctx.render(cam)
        before = ctx.read_image().copy()
        ctx.render_samples(cam, 1, 0, rt.ACCUM_REFERENCE)
        assert np.array_equal(bits(ctx.read_image()), bits(before)) and ctx.stats().samples_done == 1


def mesh_scene(parts, n_spheres=6, seed=3):
    """Several displaced icosphere meshes (each its own model) + a few spheres + floor, in the reference's host layouts.
    parts = [(subdiv, centre, radius, material)]; the meshes are star shaped, NOT convex, so the reference's
    first-front-facing rule (cl:291-303) and the nearest rule give different pictures."""
    from rtb200 import synth
    arrays, _ = synth.scene(n_spheres=n_spheres, mesh_subdiv=None, seed=seed, sphere_box=((-4, -2, 4), (4, 1, 12)), radii=(0.2, 0.6))
    verts, uvs, idx, meshes, models = [], [], [], [], []
    nv = ni = 0
    for subdiv, centre, radius, mat in parts:
        v, f = synth.icosphere(subdiv)
        bump = 1.0 + 0.35 * np.sin(5 * v[:, 0] + 1.3) * np.sin(4 * v[:, 1] + 0.4) * np.sin(6 * v[:, 2] + 2.1)
        v = v * bump[:, None] * radius + np.asarray(centre)
        v4 = np.zeros((len(v), 4), np.float32); v4[:, :3] = v
        verts.append(v4); uvs.append(np.zeros((len(v), 2), np.float32)); idx.append(f.astype(np.uint32).reshape(-1))
        meshes.append([nv, ni, len(f), 0xFFFFFFFF]); models.append([len(models), 1, mat])
        nv += len(v); ni += 3 * len(f)
    arrays.update(vertices=synth._bytes(np.concatenate(verts)), uvs=synth._bytes(np.concatenate(uvs)),
                  indices=synth._bytes(np.concatenate(idx).reshape(-1, 1)), meshes=synth._bytes(np.array(meshes, np.uint32)),
                  models=synth._bytes(np.array(models, np.uint32)))
    return arrays


@pytest.mark.parametrize("mesh_mode", [0, 1])
def test_long_meshes_in_the_brute_force_path(port, mesh_mode):
    """80- and 320-triangle meshes (3 and 10 chunks of 32 in the warp-cooperative stage) under both mesh rules."""
    import rtb200 as rt
    arrays = mesh_scene([(1, (-1.6, -0.4, 7.0), 1.3, 3), (2, (1.7, -0.2, 8.0), 1.5, 1)])
    got, want, st, rays_want = render_both(port, arrays, 200, 120, 3, 8, mesh_mode=mesh_mode, accel=rt.ACCEL_BRUTE)
    assert st.bvh_nodes == 0
    assert int((bits(got) != bits(want)).any(-1).sum()) <= 1 and abs(int(st.rays_last) - int(rays_want)) <= 20
    tri = (want[..., :3] > 0).any()
    assert tri


def test_mesh_rules_differ_on_non_convex_meshes(port):
    from rtb200 import fixtures as fx
    arrays = mesh_scene([(2, (0.0, -0.3, 6.0), 2.0, 3)], n_spheres=0)
    d = op.make_desc(arrays)
    cam = fx.camera(60, 1.5)
    a, _ = port.primary_ids(d, cam, 150, 100, op.MESH_FIRST_FRONT)
    b, _ = port.primary_ids(d, cam, 150, 100, op.MESH_NEAREST)
    assert (a != b).mean() > 0.002                       # the test above really exercises two different rules


def test_many_mesh_instances(port):
    """Five mesh instances (each with its own conservative bound) + spheres: strict image == oracle."""
    import rtb200 as rt
    parts = [(0, (-3.0 + 1.5 * k, -0.5 + 0.2 * k, 6.0 + 0.7 * k), 0.7, 1 + k % 6) for k in range(5)]
    arrays = mesh_scene(parts, n_spheres=10)
    got, want, st, rays_want = render_both(port, arrays, 160, 100, 4, 10, mesh_mode=0, accel=rt.ACCEL_BRUTE)
    assert int((bits(got) != bits(want)).any(-1).sum()) <= 1 and abs(int(st.rays_last) - int(rays_want)) <= 20


@pytest.mark.parametrize("arith", ["strict", "fast"])
def test_largest_brute_force_scene(port, arith):
    """Maximum size of the shared-memory path: ~9 900 spheres fill the 160 KB scene budget; the render kernels add their
    per-lane state on top and must still launch.  Strict image == oracle; the fast arm (sample chunks) counts the same
    number of segments per sample within Monte-Carlo noise."""
    import rtb200 as rt
    from rtb200 import fixtures as fx, synth
    arrays, _ = synth.scene(n_spheres=9900, mesh_subdiv=None, seed=4, sphere_box=((-10, -5, 4), (10, 1.2, 30)), radii=(0.05, 0.15))
    W, H, depth = 48, 24, 4
    cam = fx.camera(60, W / H)
    if arith == "strict":
        got, want, st, rays_want = render_both(port, arrays, W, H, 2, depth, accel=rt.ACCEL_BRUTE)
        assert st.bvh_nodes == 0
        assert np.array_equal(bits(got), bits(want)) and st.rays_last == rays_want
        return
    with rt.Context(W, H) as ctx:
        ctx.set_params(max_depth=depth, rng_mode=rt.RNG_COUNTER, seed=13, arith=rt.ARITH_FAST, accel=rt.ACCEL_BRUTE)
        ctx.upload_scene(arrays)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, 64, rt.ACCUM_SUM)
        acc, st = ctx.read_accum(), ctx.stats()
        assert st.bvh_nodes == 0 and np.all(acc[..., 3] == 64) and np.isfinite(acc).all()
        ctx.set_params(max_depth=depth, rng_mode=rt.RNG_COUNTER, seed=13, arith=rt.ARITH_STRICT, accel=rt.ACCEL_BRUTE)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, 64, rt.ACCUM_SUM)
        ref, st_ref = ctx.read_accum(), ctx.stats()
    assert abs(int(st.rays_last) - int(st_ref.rays_last)) <= 0.02 * st_ref.rays_last
    assert abs(float(acc[..., :3].mean()) - float(ref[..., :3].mean())) <= 0.05 * max(1.0, float(ref[..., :3].mean()))
