"""rt_trace_rays_noalloc (include/rt_capi.h): the nearest-hit probe with its device scratch in a static device array instead of
per-call cudaMalloc — for callers that may not allocate.  No counterpart in the reference.
Checked through the C ABI against the existing entry point rt_trace_rays in strict arithmetic, bit for bit (the
bar of the other probe tests), on consecutive calls that share the static scratch, on the brute-force path and through the
LBVH.  This file is collected last on purpose: it must not hide other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def random_rays(n, seed, box=((-14, -8, -2), (14, 3, 45))):
    rng = np.random.RandomState(seed)
    o = rng.uniform(box[0], box[1], (n, 3)).astype(np.float32)
    v = rng.normal(size=(n, 3))
    v = (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)
    return np.concatenate([o, v], 1)


def test_noalloc_probe_equals_trace_rays_on_consecutive_calls(scenes, table):
    import rtb200 as rt
    arr, tex = scenes["default"]
    with rt.Context(64, 64) as ctx:
        ctx.upload_textures(tex)
        ctx.upload_scene(arr)
        ctx.set_random_table(table)
        ctx.set_params(arith=rt.ARITH_STRICT)
        # two calls in a row share the static array: a large batch first, then a smaller, different one (nothing stale may leak)
        for n, seed in ((rt.NOALLOC_MAX_RAYS, 1), (12345, 2), (1, 3)):
            rays = random_rays(n, seed, box=((-3, -2, -1), (3, 2, 6)))
            want_id, want_t = ctx.trace_rays(rays)
            got_id, got_t = ctx.trace_rays_noalloc(rays)
            assert (want_id != rt.ID_MISS).mean() > 0.3
            assert np.array_equal(got_id, want_id) and np.array_equal(bits(got_t), bits(want_t))
        # the result is the strict-arithmetic one whatever the context's arithmetic is
        rays = random_rays(4000, 4, box=((-3, -2, -1), (3, 2, 6)))
        want_id, want_t = ctx.trace_rays(rays)
        ctx.set_params(arith=rt.ARITH_FAST)
        got_id, got_t = ctx.trace_rays_noalloc(rays)
        assert np.array_equal(got_id, want_id) and np.array_equal(bits(got_t), bits(want_t))
        # beyond the static scratch: a clear error, nothing launched
        with pytest.raises(rt.RtError) as e:
            ctx.trace_rays_noalloc(random_rays(rt.NOALLOC_MAX_RAYS + 1, 5))
        assert e.value.code == 1 and "exceed the static scratch" in str(e.value)
        ids, _ = ctx.trace_rays_noalloc(np.zeros((0, 6), np.float32))
        assert ids.size == 0


def test_noalloc_probe_through_the_lbvh_builds_nothing():
    """Large scenes: the entry point does not build the LBVH (that allocates) — it reports RT_ERR_STATE until another call has
    built it, then answers through it exactly like rt_trace_rays, twice in a row."""
    import rtb200 as rt
    from rtb200 import synth
    arrays, info = synth.scene(n_spheres=800, mesh_subdiv=4, seed=3)
    with rt.Context(64, 64) as ctx:
        ctx.set_params(arith=rt.ARITH_STRICT, mesh_mode=rt.MESH_NEAREST)
        ctx.upload_scene(arrays)
        rays = random_rays(30000, 7)
        with pytest.raises(rt.RtError) as e:
            ctx.trace_rays_noalloc(rays)
        assert e.value.code == 3 and "LBVH not built" in str(e.value)
        want_id, want_t = ctx.trace_rays(rays)                  # builds the LBVH
        assert ctx.stats().bvh_nodes == info["spheres"] + info["triangles"] - 1
        for _ in range(2):
            got_id, got_t = ctx.trace_rays_noalloc(rays)
            assert np.array_equal(got_id, want_id) and np.array_equal(bits(got_t), bits(want_t))
        rays2 = random_rays(5000, 8)
        want2 = ctx.trace_rays(rays2)
        got2 = ctx.trace_rays_noalloc(rays2)
        assert np.array_equal(got2[0], want2[0]) and np.array_equal(bits(got2[1]), bits(want2[1]))
        assert len(np.unique(want_id >> 28)) >= 3
