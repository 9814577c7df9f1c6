"""GPU tests of the round-2 additions to the C ABI: deferred render()/renderAgain() batching (bit-identical to the eager
schedule of the reference), render regions, the counted-work (instrumented) LBVH kernel, the head-class ordering of the
stochastic list, and the error path of a failed LBVH build."""
import numpy as np
import pytest

import oracle_py as op

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def make_ctx(rt, scenes, name, W, H, table, **params):
    arr, tex = scenes[name]
    ctx = rt.Context(W, H)
    ctx.upload_textures(tex)
    ctx.upload_scene(arr)
    ctx.set_random_table(table)
    ctx.set_params(**params)
    return ctx


@pytest.mark.parametrize("arith", ["strict", "fast"])
def test_deferred_render_again_equals_eager_schedule(scenes, table, monkeypatch, arith):
    """render() + renderAgain() x (n-1) recorded and rendered by one launch (the default) == one launch and one host round
    trip per sample, the reference's schedule (src/raytracer.cpp:137-140,158-161; RTB200_EAGER=1): same image bits, same
    segment count — also across a camera change, a statistics read in the middle and a restart with render()."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    W, H = 320, 200
    cam_a, cam_b = fx.camera(60, W / H), fx.camera(50, W / H, (0.2, 0.0, -0.3), 10.0, -5.0)
    res = {}
    for mode in ("deferred", "eager"):
        if mode == "eager":
            monkeypatch.setenv("RTB200_EAGER", "1")
        else:
            monkeypatch.delenv("RTB200_EAGER", raising=False)
        with make_ctx(rt, scenes, "default", W, H, table, max_depth=12, arith=getattr(rt, "ARITH_" + arith.upper())) as ctx:
            ctx.render(cam_a)
            for _ in range(4):
                ctx.render_again(cam_a)
            ctx.render(cam_b)                       # camera moved: the five samples above are discarded unseen
            for _ in range(6):
                ctx.render_again(cam_b)
            mid = ctx.stats()                       # observation: flushes 7 samples
            assert mid.samples_done == 7
            for _ in range(5):
                ctx.render_again(cam_b)
            img = ctx.read_image().copy()
            st = ctx.stats()
            launches = st.kernel_launches
            res[mode] = (img, st.rays_total, st.samples_done, mid.rays_total, launches)
    a, b = res["deferred"], res["eager"]
    assert np.array_equal(bits(a[0]), bits(b[0]))
    assert a[1] == b[1] and a[2] == b[2] == 12 and a[3] == b[3]
    assert a[4] < b[4]                              # fewer launches for the same picture


def same_sums(a, b, exact):
    """Accumulation buffers of the same samples: bit for bit, or (fast arm with sample stealing, TraceLaunch::steal: the samples
    of a pixel are added in an order that depends on the other pixels of its warp) equal to rounding, sample counts exact."""
    if exact:
        return np.array_equal(bits(a), bits(b))
    return np.array_equal(a[..., 3], b[..., 3]) and np.allclose(a[..., :3], b[..., :3], rtol=1e-5, atol=1e-5)


def test_sample_stealing_is_reproducible_and_keeps_the_sums(scenes, table, monkeypatch):
    """TraceLaunch::steal (k_trace_tail_fast, linear sums): lanes out of samples take samples of the busiest lane of their warp.
    The same call gives the same bits every time; against the launch without stealing the samples are the same ones (sample
    counts, rays and queries equal) added in another order (sums equal to rounding)."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    W, H, spp = 320, 200, 48
    cam = fx.camera(60, W / H)
    res = {}
    for key, env in (("on_a", "1"), ("on_b", "1"), ("off", "0")):
        monkeypatch.setenv("RTB200_TAIL_STEAL", env)
        with make_ctx(rt, scenes, "default", W, H, table, max_depth=12, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=9) as ctx:
            ctx.clear_accum()
            ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
            st = ctx.stats()
            res[key] = (ctx.read_accum().copy(), st.rays_last, st.queries_last)
    assert np.array_equal(bits(res["on_a"][0]), bits(res["on_b"][0]))
    assert res["on_a"][1:] == res["on_b"][1:] == res["off"][1:]
    assert same_sums(res["on_a"][0], res["off"][0], exact=False)
    assert not np.array_equal(bits(res["on_a"][0]), bits(res["off"][0]))       # stealing did happen
    assert np.all(res["on_a"][0][..., 3] == spp)


def test_region_renders_the_same_pixels_as_the_full_frame(scenes, table, monkeypatch):
    """rt_set_region: the work-items of a window are those of the full frame (same primary rays, same RNG keys), everything
    outside keeps its contents; bad windows are refused.  Strict arm, and fast arm without sample stealing: bit for bit."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    W, H, spp = 400, 240, 6
    cam = fx.camera(60, W / H)
    x0, y0, w, h = 123, 57, 90, 101
    for arith, stealing in ((rt.ARITH_STRICT, True), (rt.ARITH_FAST, True), (rt.ARITH_FAST, False)):
        monkeypatch.setenv("RTB200_TAIL_STEAL", "1" if stealing else "0")
        exact = arith == rt.ARITH_STRICT or not stealing
        for head in (rt.HEAD_ON, rt.HEAD_OFF):
            with make_ctx(rt, scenes, "default", W, H, table, max_depth=10, arith=arith, rng_mode=rt.RNG_COUNTER_TABLE, seed=3, path_head=head) as ctx:
                ctx.clear_accum()
                ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
                full = ctx.read_accum().copy()
                rays_full = ctx.stats().rays_last
                ctx.clear_accum()
                ctx.set_region(x0, y0, w, h)
                ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
                part = ctx.read_accum().copy()
                rays_part = ctx.stats().rays_last
                assert same_sums(part[y0:y0 + h, x0:x0 + w], full[y0:y0 + h, x0:x0 + w], exact)
                part[y0:y0 + h, x0:x0 + w] = 0
                assert not part.any() and 0 < rays_part < rays_full
                ctx.set_region()                                   # back to the whole frame
                ctx.clear_accum()
                ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
                assert np.array_equal(bits(ctx.read_accum()), bits(full))
                for bad in ((-1, 0, 10, 10), (0, 0, W + 1, 10), (W - 5, 0, 6, 10), (0, H - 1, 4, 2)):
                    with pytest.raises(rt.RtError):
                        ctx.set_region(*bad)


def test_head_class_ordering_changes_nothing(scenes, table, monkeypatch):
    """The stochastic list ordered by (material type, primitive kind, primitive) of the head hit instead of by tile: only the
    assignment of pixels to lanes changes — strict images, segment and query counts are bit-identical."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    W, H, spp = 480, 270, 40
    cam = fx.camera(60, W / H)
    res = {}
    for name in ("default", "c3"):
        for mask in ("0x3FFF", "0"):
            monkeypatch.setenv("RTB200_HEAD_SORT", mask)
            with make_ctx(rt, scenes, name, W, H, table, max_depth=30, arith=rt.ARITH_STRICT, rng_mode=rt.RNG_COUNTER_TABLE, seed=2) as ctx:
                ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
                st = ctx.stats()
                res[(name, mask)] = (ctx.read_image().copy(), st.rays_last, st.queries_last)
        a, b = res[(name, "0x3FFF")], res[(name, "0")]
        assert np.array_equal(bits(a[0]), bits(b[0])) and a[1:] == b[1:]


def test_counted_work_of_the_lbvh_kernel(table):
    """rt_set_work_counting: the instrumented copy of the LBVH render kernel renders the SAME image and reports plausible
    counts (every query visits nodes, tests leaves of the kinds the scene has, scans the plane list)."""
    import rtb200 as rt
    from rtb200 import fixtures as fx, synth
    arrays, info = synth.scene(n_spheres=3000, mesh_subdiv=4, seed=5)
    W, H, spp = 320, 180, 8
    cam = fx.camera(60, W / H)
    with rt.Context(W, H) as ctx:
        ctx.upload_scene(arrays)
        ctx.set_random_table(table)
        ctx.set_params(max_depth=12, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, arith=rt.ARITH_FAST, mesh_mode=rt.MESH_NEAREST)
        ctx.render_samples(cam, 0, 1, rt.ACCUM_SUM)            # builds the path heads (their queries count once)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        plain, q_plain = ctx.read_accum().copy(), ctx.stats().queries_last
        ctx.set_work_counting(True)
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        counted, q_counted = ctx.read_accum().copy(), ctx.stats().queries_last
        wc = ctx.work_counts()
        ctx.set_work_counting(False)
    assert ctx is not None and np.array_equal(bits(plain), bits(counted)) and q_plain == q_counted
    assert wc["queries"] == q_counted > W * H
    assert wc["plane_tests"] == wc["queries"] * info["planes"]
    assert wc["node_visits"] > 3 * wc["queries"] and wc["sphere_tests"] > 0 and wc["triangle_tests"] > 0 and wc["lens_tests"] == 0
    assert wc["node_visits"] < 400 * wc["queries"]


def test_failed_lbvh_build_is_an_error_not_a_half_built_tree(table, monkeypatch):
    """a build that fails half way must not leave a tree that the next render traverses.  The failure is injected
    (RTB200_TEST_FAIL_BVH_BUILD); the render call reports RT_ERR_CUDA, and once the cause is gone the next call rebuilds."""
    import rtb200 as rt
    from rtb200 import fixtures as fx, synth
    arrays, _ = synth.scene(n_spheres=1500, mesh_subdiv=None, seed=9)
    W, H = 160, 96
    cam = fx.camera(60, W / H)
    with rt.Context(W, H) as ctx:
        ctx.upload_scene(arrays)
        ctx.set_random_table(table)
        ctx.set_params(max_depth=8, rng_mode=rt.RNG_COUNTER_TABLE, seed=1, arith=rt.ARITH_STRICT, mesh_mode=rt.MESH_NEAREST)
        monkeypatch.setenv("RTB200_TEST_FAIL_BVH_BUILD", "1")
        for _ in range(2):
            with pytest.raises(rt.RtError) as e:
                ctx.render_samples(cam, 0, 2, rt.ACCUM_REFERENCE)
            assert e.value.code == 2 and "build_lbvh" in str(e.value)
            assert ctx.stats().bvh_nodes == 0
        monkeypatch.delenv("RTB200_TEST_FAIL_BVH_BUILD")
        ctx.render_samples(cam, 0, 2, rt.ACCUM_REFERENCE)
        assert ctx.stats().bvh_nodes == 1500 and np.isfinite(ctx.read_image()).all()


def test_pixel_split_parts_add_up_bit_for_bit(scenes, table, monkeypatch):
    """rt_set_pixel_split: the parts of a frame (interleaved 128-pixel groups, all samples each) rendered one after the
    other are disjoint, cover the frame, and their sum IS the single-context accumulation buffer — in both arms, with path
    heads on and off, brute force and LBVH (fast arm, brute force: bit for bit without sample stealing, to rounding with it)."""
    import rtb200 as rt
    from rtb200 import fixtures as fx
    W, H, spp = 400, 236, 10
    cam = fx.camera(60, W / H)
    for arith, head, accel, mesh, stealing in ((rt.ARITH_FAST, rt.HEAD_ON, rt.ACCEL_AUTO, rt.MESH_FIRST_FRONT, False),
                                               (rt.ARITH_FAST, rt.HEAD_ON, rt.ACCEL_AUTO, rt.MESH_FIRST_FRONT, True),
                                               (rt.ARITH_STRICT, rt.HEAD_OFF, rt.ACCEL_AUTO, rt.MESH_FIRST_FRONT, True),
                                               (rt.ARITH_FAST, rt.HEAD_ON, rt.ACCEL_BVH, rt.MESH_NEAREST, True),
                                               (rt.ARITH_STRICT, rt.HEAD_OFF, rt.ACCEL_BVH, rt.MESH_NEAREST, True)):
        monkeypatch.setenv("RTB200_TAIL_STEAL", "1" if stealing else "0")
        exact = not (stealing and arith == rt.ARITH_FAST and accel == rt.ACCEL_AUTO)     # the LBVH kernel has no stealing
        with make_ctx(rt, scenes, "default", W, H, table, max_depth=12, arith=arith, rng_mode=rt.RNG_COUNTER_TABLE, seed=6, path_head=head,
                      accel=accel, mesh_mode=mesh) as ctx:
            ctx.clear_accum()
            ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
            whole, rays_whole = ctx.read_accum().copy(), ctx.stats().rays_last
            parts, rays = np.zeros_like(whole), 0
            for n_parts in (3,):
                for part in range(n_parts):
                    ctx.set_pixel_split(n_parts, part)
                    ctx.clear_accum()
                    ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
                    a = ctx.read_accum()
                    assert not (parts[..., 3] * a[..., 3]).any()            # disjoint
                    parts += a
                    rays += ctx.stats().rays_last
            assert same_sums(parts, whole, exact) and rays == rays_whole
            with pytest.raises(rt.RtError):
                ctx.set_pixel_split(4, 4)
            ctx.set_pixel_split(1, 0)


def test_finalize_target_in_mapped_host_memory(scenes, table):
    """rt_set_finalize_target: rt_finalize stores the picture straight into page-locked host memory (no staging copy) and it equals
    the image read back the ordinary way; pageable memory is refused."""
    import torch
    import rtb200 as rt
    from rtb200 import fixtures as fx
    W, H, spp = 320, 180, 8
    cam = fx.camera(60, W / H)
    host = torch.empty((H, W, 4), dtype=torch.float32).pin_memory()
    host.zero_()
    with make_ctx(rt, scenes, "default", W, H, table, max_depth=10, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=4) as ctx:
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        ctx.finalize(spp)
        want = ctx.read_image().copy()
        ctx.set_finalize_target(host.data_ptr(), host.numel() * 4)
        ctx.finalize(spp)                                   # synchronous: the stores have landed
        assert np.array_equal(bits(host.numpy()), bits(want))
        pageable = np.zeros((H, W, 4), np.float32)
        with pytest.raises(rt.RtError):
            ctx.set_finalize_target(pageable.ctypes.data, pageable.nbytes)
        with pytest.raises(rt.RtError):
            ctx.set_finalize_target(host.data_ptr(), 16)
        ctx.set_finalize_target(None, 0)
