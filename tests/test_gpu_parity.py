"""GPU parity tests (run on the B200 box: pytest -m gpu).  Everything goes through the C ABI (rtb200 is a
ctypes binding of include/rt_capi.h) and is checked against the oracle on the same seeded inputs.

Bars (stated here, repeated in DESIGN.md):
  STRICT arithmetic  bit-exact: image bits, ray counts, primary ids and t, single-bounce outputs.
                     (A dielectric pow() rounding flip is possible with probability ~1e-8 per bounce; the
                      image tests allow <= 1 differing pixel per frame for that single reason and print the observed count.)
  FAST arithmetic    identical primary-hit ids; single-bounce outputs within 2e-4 absolute for
                     non-chaotic (non-table-hashed) results; image RMSE(gamma) vs oracle <= 0.03 at 16 spp
                     (the reference's own FMA-on/off difference is 0.022) and mean difference <= 2e-3.
  COUNTER RNG        strict arithmetic is bit-exact against the oracle's counter mode; statistically equal
                     to table mode: 16x16 block-mean RMSE <= 2x the table seed-to-seed value.
"""
import numpy as np
import pytest

import oracle_py as op

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def rt():
    import rtb200
    return rtb200


def make_ctx(rt, scenes, name, W, H, table, **params):
    arr, tex = scenes[name]
    ctx = rt.Context(W, H)
    ctx.upload_textures(tex)
    ctx.upload_scene(arr)
    ctx.set_random_table(table)
    ctx.set_params(**params)
    return ctx


def n_diff_pixels(a, b):
    """Pixels whose bits differ; the observed count is printed (pytest -s / the captured output of a failure) so that the
    allowance of ONE pixel per frame — a powf double-rounding flip of the Schlick term, p ~ 1e-8 per dielectric bounce —
    cannot hide anything: the expected count at 480 k pixels x 16 spp is 0.05."""
    n = int((bits(a[..., :3]) != bits(b[..., :3])).any(-1).sum())
    print("differing pixels: %d of %d" % (n, a.shape[0] * a.shape[1]))
    return n


@pytest.mark.parametrize("name,W,H,spp,depth", [("default", 800, 600, 16, 8), ("default", 320, 200, 8, 30),
                                               ("c3", 480, 270, 8, 30), ("default", 101, 67, 3, 2)])
def test_strict_image_bit_exact(rt, port, scenes, table, name, W, H, spp, depth):
    """BASELINE config C1 (800x600, 16 spp, depth 8) and friends: strict CUDA == oracle, bit for bit."""
    from rtb200 import fixtures as fx
    arr, tex = scenes[name]
    cam = fx.camera(60, W / H)
    want, rays_want, _ = port.render(op.make_desc(arr), tex, cam, table, W, H, 0, spp, port.params(max_depth=depth))
    with make_ctx(rt, scenes, name, W, H, table, max_depth=depth, arith=rt.ARITH_STRICT) as ctx:
        # (a) the reference's own call sequence: render() then renderAgain() x (spp-1)
        ctx.render(cam)
        rays = ctx.stats().rays_last
        for _ in range(spp - 1):
            ctx.render_again(cam)
            rays += ctx.stats().rays_last
        got = ctx.read_image()
        assert n_diff_pixels(got, want) <= 1
        assert abs(int(rays) - int(rays_want)) <= 60
        assert ctx.stats().rays_total == rays and ctx.stats().samples_done == spp
        assert np.all(got[..., 3] == 1.0)
        # (b) all samples in one launch: identical bits to the sequential calls
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        got2 = ctx.read_image()
        assert np.array_equal(bits(got2), bits(got))
        assert ctx.stats().rays_last == rays
        # (c) split launches continue the chain
        ctx.render_samples(cam, 0, spp // 2, rt.ACCUM_REFERENCE)
        ctx.render_samples(cam, spp // 2, spp - spp // 2, rt.ACCUM_REFERENCE)
        assert np.array_equal(bits(ctx.read_image()), bits(got))


@pytest.mark.parametrize("name", ["default", "c3"])
def test_golden_fixture_images(rt, scenes, golden, table, name):
    """Against the committed outputs of the reference kernel build (no oracle code involved at run time)."""
    g = golden[name]
    W, H, cam = int(g["W"]), int(g["H"]), g["cam"]
    with make_ctx(rt, scenes, name, W, H, table, max_depth=30, arith=rt.ARITH_STRICT) as ctx:
        ctx.render(cam)
        assert n_diff_pixels(ctx.read_image(), g["image_1spp"]) <= 1
        assert abs(int(ctx.stats().rays_last) - int(g["rays_1spp"])) <= 20
        for _ in range(3):
            ctx.render_again(cam)
        assert n_diff_pixels(ctx.read_image(), g["image_4spp"]) <= 1
        ids, ts = ctx.primary_ids(cam)
        assert np.array_equal(ids, g["ids"]) and np.array_equal(bits(ts), bits(g["ts"]))
        ctx.set_params(max_depth=8, arith=rt.ARITH_STRICT)
        ctx.render_samples(cam, 0, 4, rt.ACCUM_REFERENCE)
        assert n_diff_pixels(ctx.read_image(), g["image_4spp_depth8"]) <= 1


@pytest.mark.parametrize("arith", ["strict", "fast"])
@pytest.mark.parametrize("name,W,H", [("default", 800, 600), ("default", 1920, 1080), ("c3", 1920, 1080), ("c3", 480, 270)])
def test_primary_ids_identical(rt, port, scenes, golden, table, name, W, H, arith):
    from rtb200 import fixtures as fx
    arr, _ = scenes[name]
    is_golden_size = (W, H) == (int(golden[name]["big_W"]), int(golden[name]["big_H"]))
    cam = golden[name]["big_cam"] if is_golden_size else fx.camera(60, W / H)
    want_ids, want_t = port.primary_ids(op.make_desc(arr), cam, W, H)
    with make_ctx(rt, scenes, name, W, H, table, arith=getattr(rt, "ARITH_" + arith.upper())) as ctx:
        ids, ts = ctx.primary_ids(cam)
    # fast arithmetic certifies its primary hits against the Exact policy, so ids are identical in both arms
    assert int((ids != want_ids).sum()) == 0
    if arith == "strict":
        assert np.array_equal(bits(ts), bits(want_t))
    else:
        same = ids == want_ids
        assert np.abs(ts - want_t)[same].max() <= 0.05
    if is_golden_size:
        assert np.array_equal(ids, golden[name]["big_ids"])


@pytest.mark.parametrize("name", ["default", "c3"])
def test_single_bounce_vectors(rt, scenes, golden, table, name):
    """Recorded (ray, seed, pixel) -> (new ray, throughput) tuples of the reference kernel: immune to chaos."""
    g = golden[name]
    meta = g["bounce_meta"]
    with make_ctx(rt, scenes, name, int(g["W"]), int(g["H"]), table, arith=rt.ARITH_STRICT) as ctx:
        r, col, ended = ctx.bounce_rays(g["bounce_in"], meta[:, 0] + meta[:, 1], 0, meta[:, 2], meta[:, 3])
        assert np.array_equal(bits(r), bits(g["bounce_out"]))
        assert np.array_equal(bits(col), bits(g["bounce_col"]))
        ctx.set_params(arith=rt.ARITH_FAST)
        r2, col2, _ = ctx.bounce_rays(g["bounce_in"], meta[:, 0] + meta[:, 1], 0, meta[:, 2], meta[:, 3])
    # fast arithmetic: origins (hit points) agree closely; directions of diffuse bounces depend on the table
    # hash of the direction (chaotic), so compare those only where the strict direction was reproduced closely
    assert np.abs(r2[:, :3] - g["bounce_out"][:, :3]).max() < 2e-3
    close = np.abs(r2[:, 3:] - g["bounce_out"][:, 3:]).max(1) < 1e-3
    assert close.mean() > 0.9
    assert np.abs(col2[close] - g["bounce_col"][close]).max() < 1e-3


def test_fast_image_statistics(rt, port, scenes, table):
    """C1: fast arithmetic vs oracle at 16 spp: RMSE(gamma) <= 0.03, mean difference <= 2e-3, rays within 0.1 %."""
    from rtb200 import fixtures as fx
    W, H, spp, depth = 800, 600, 16, 8
    arr, tex = scenes["default"]
    cam = fx.camera(60, W / H)
    want, rays_want, _ = port.render(op.make_desc(arr), tex, cam, table, W, H, 0, spp, port.params(max_depth=depth))
    with make_ctx(rt, scenes, "default", W, H, table, max_depth=depth, arith=rt.ARITH_FAST) as ctx:
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        got = ctx.read_image()
        rays = ctx.stats().rays_last
        # linear-sum path gives the same picture up to float summation order
        ctx.clear_accum()
        ctx.render_samples(cam, 0, spp, rt.ACCUM_SUM)
        ctx.finalize(spp)
        got_sum = ctx.read_image()
    d = got[..., :3] - want[..., :3]
    assert np.sqrt((d * d).mean()) <= 0.03
    assert abs(float(got[..., :3].mean()) - float(want[..., :3].mean())) <= 2e-3
    assert abs(int(rays) - int(rays_want)) <= 1e-3 * rays_want
    assert np.abs(got_sum[..., :3] - got[..., :3]).max() < 1e-4


def block_means(img, b=16):
    H, W = img.shape[:2]
    lin = img[: H // b * b, : W // b * b, :3].astype(np.float64) ** 2      # linear radiance
    return lin.reshape(H // b, b, W // b, b, 3).mean((1, 3))


def test_counter_rng_strict_bit_exact_and_same_law(rt, port, scenes, table):
    from rtb200 import fixtures as fx
    W, H, spp = 320, 240, 64
    arr, tex = scenes["default"]
    d = op.make_desc(arr)
    cam = fx.camera(60, W / H)
    want, rays_want, _ = port.render(d, tex, cam, None, W, H, 0, spp, port.params(max_depth=30, rng_mode=op.RNG_COUNTER, seed=77))
    with make_ctx(rt, scenes, "default", W, H, table, max_depth=30, arith=rt.ARITH_STRICT, rng_mode=rt.RNG_COUNTER, seed=77) as ctx:
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        got = ctx.read_image()
        assert n_diff_pixels(got, want) <= 1
        assert abs(int(ctx.stats().rays_last) - int(rays_want)) <= 60
        ctx.set_params(max_depth=30, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER, seed=78)
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        fast_c = ctx.read_image()
    # the law: counter-mode block means agree with table-mode block means as well as two table seeds agree
    t_a, _, _ = port.render(d, tex, cam, table, W, H, 0, spp, port.params(max_depth=30))
    t_b, _, _ = port.render(d, tex, cam, port.make_table(999), W, H, 0, spp, port.params(max_depth=30))
    noise = np.sqrt(((block_means(t_a) - block_means(t_b)) ** 2).mean())
    for img in (got, fast_c):
        err = np.sqrt(((block_means(img) - block_means(t_a)) ** 2).mean())
        assert err <= 2.0 * noise, (err, noise)
        assert abs(float((img[..., :3] ** 2).mean()) - float((t_a[..., :3] ** 2).mean())) <= 0.004


def test_counter_table_rng_strict_bit_exact_and_same_law(rt, port, scenes, table):
    """RT_RNG_COUNTER_TABLE (bench.py's generator): the reference's table entered at a counter-keyed index.  Strict
    arithmetic is bit-exact against the oracle's restatement; block means agree with the reference's table mode as
    well as two table seeds agree with each other."""
    from rtb200 import fixtures as fx
    W, H, spp = 320, 240, 64
    arr, tex = scenes["default"]
    d = op.make_desc(arr)
    cam = fx.camera(60, W / H)
    want, rays_want, _ = port.render(d, tex, cam, table, W, H, 0, spp, port.params(max_depth=30, rng_mode=op.RNG_COUNTER_TABLE, seed=77))
    with make_ctx(rt, scenes, "default", W, H, table, max_depth=30, arith=rt.ARITH_STRICT, rng_mode=rt.RNG_COUNTER_TABLE, seed=77) as ctx:
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        got = ctx.read_image()
        assert n_diff_pixels(got, want) <= 1
        assert abs(int(ctx.stats().rays_last) - int(rays_want)) <= 60
        ctx.set_params(max_depth=30, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=78)
        ctx.render_samples(cam, 0, spp, rt.ACCUM_REFERENCE)
        fast_c = ctx.read_image()
    t_a, _, _ = port.render(d, tex, cam, table, W, H, 0, spp, port.params(max_depth=30))
    t_b, _, _ = port.render(d, tex, cam, port.make_table(999), W, H, 0, spp, port.params(max_depth=30))
    noise = np.sqrt(((block_means(t_a) - block_means(t_b)) ** 2).mean())
    for img in (got, fast_c):
        err = np.sqrt(((block_means(img) - block_means(t_a)) ** 2).mean())
        assert err <= 2.0 * noise, (err, noise)
        assert abs(float((img[..., :3] ** 2).mean()) - float((t_a[..., :3] ** 2).mean())) <= 0.004


def test_sample_split_sums_to_whole(rt, scenes, table):
    """What the multi-GPU split relies on: disjoint global sample ranges accumulate to the single-launch sum."""
    from rtb200 import fixtures as fx
    W, H = 256, 128
    cam = fx.camera(60, W / H)
    with make_ctx(rt, scenes, "default", W, H, table, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER, seed=5) as ctx:
        ctx.clear_accum()
        ctx.render_samples(cam, 0, 12, rt.ACCUM_SUM)
        whole = ctx.read_accum()
        parts = np.zeros_like(whole)
        for first, n in ((0, 3), (3, 3), (6, 6)):
            ctx.clear_accum()
            ctx.render_samples(cam, first, n, rt.ACCUM_SUM)
            parts += ctx.read_accum()
    assert np.all(whole[..., 3] == 12) and np.all(parts[..., 3] == 12)
    assert np.abs(parts - whole).max() <= 1e-5 * max(1.0, float(np.abs(whole).max()))


def test_sample_chunks_change_only_the_summation_order(rt, scenes, table, monkeypatch):
    """The fast arm's tail kernel splits the samples of a pixel into chunks (work items small enough that the last wave
    of blocks does not idle the SMs) and adds the chunk sums in chunk order: the paths are the same — the RNG is keyed
    by the global sample index — so the segment and query counts are identical and the sums differ by rounding only."""
    from rtb200 import fixtures as fx
    W, H, spp = 320, 180, 200
    cam = fx.camera(60, W / H)
    res = {}
    with make_ctx(rt, scenes, "default", W, H, table, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=9) as ctx:
        ctx.render_samples(cam, 0, 1, rt.ACCUM_SUM)           # builds the path heads (their queries count once)
        for chunks in (1, 2, 7, 12, 0):                      # 7: ragged last chunk; 12: capped at spp / 20; 0: heuristic
            if chunks:
                monkeypatch.setenv("RTB200_TAIL_CHUNKS", str(chunks))
            else:
                monkeypatch.delenv("RTB200_TAIL_CHUNKS")
            ctx.clear_accum()
            ctx.render_samples(cam, 3, spp, rt.ACCUM_SUM)
            st = ctx.stats()
            res[chunks] = (ctx.read_accum().copy(), st.rays_last, st.queries_last)
            ctx.render_samples(cam, 3, 40, rt.ACCUM_SUM)     # accumulates on top of what is there
            assert np.all(ctx.read_accum()[..., 3] == spp + 40)
    base = res[1]
    assert base[0][..., :3].max() > 10.0
    for chunks, (acc, rays, queries) in res.items():
        assert np.all(acc[..., 3] == spp)
        assert rays == base[1] and queries == base[2]
        assert np.abs(acc - base[0]).max() <= 2e-6 * float(base[0].max()) * spp ** 0.5, chunks
    assert not np.array_equal(res[7][0], base[0])            # the chunked path really ran


@pytest.mark.parametrize("arith,accum", [("strict", "reference"), ("strict", "sum"), ("fast", "sum"), ("fast", "sum-stealing")])
def test_longest_blocks_first_changes_nothing(rt, scenes, table, monkeypatch, arith, accum):
    """Frames of >= 2048 pixel blocks: the first two samples of a call are a pilot launch that counts the queries of every
    block of 128 stochastic pixels, the rest is launched with the most expensive blocks first (and the order is kept
    while the path heads are valid).  Only the dispatch order changes: images, segment and query counts are bit-identical
    to the launch in list order (fast arm: with whole pixels as work items in both, and without sample stealing — with it the
    pilot launch changes which lane adds which sample, and the sums agree to rounding)."""
    from rtb200 import fixtures as fx
    W, H, spp = 1024, 576, 40
    cam = fx.camera(60, W / H)
    A = rt.ARITH_STRICT if arith == "strict" else rt.ARITH_FAST
    mode = rt.ACCUM_REFERENCE if accum == "reference" else rt.ACCUM_SUM
    stealing = accum == "sum-stealing"
    accum = "sum" if stealing else accum
    monkeypatch.setenv("RTB200_TAIL_STEAL", "1" if stealing else "0")
    monkeypatch.setenv("RTB200_TAIL_CHUNKS", "1")
    res = {}
    for lpt in ("0", "1"):
        monkeypatch.setenv("RTB200_TAIL_LPT", lpt)
        with make_ctx(rt, scenes, "default", W, H, table, arith=A, rng_mode=rt.RNG_COUNTER_TABLE, seed=2) as ctx:
            ctx.clear_accum()
            l0 = ctx.stats().kernel_launches
            ctx.render_samples(cam, 0, spp, mode)
            st = ctx.stats()
            l1 = st.kernel_launches
            first = (ctx.read_image().copy() if accum == "reference" else ctx.read_accum().copy(), st.rays_last, st.queries_last)
            ctx.render_samples(cam, spp, spp, mode)          # continues; the order of the first call is reused
            st = ctx.stats()
            second = (ctx.read_image().copy() if accum == "reference" else ctx.read_accum().copy(), st.rays_last, st.queries_last)
            res[lpt] = (first, second, l1 - l0, st.kernel_launches - l1)
    off, on = res["0"], res["1"]
    for a, b in ((off[0], on[0]), (off[1], on[1])):
        if stealing:
            assert np.array_equal(a[0][..., 3], b[0][..., 3]) and np.allclose(a[0][..., :3], b[0][..., :3], rtol=1e-5, atol=1e-5)
        else:
            assert np.array_equal(bits(a[0]), bits(b[0]))
        assert a[1] == b[1] and a[2] == b[2]
    assert on[2] == off[2] + 3 and on[3] == off[3]           # pilot + sort (2 launches) once, then the cached order
    if accum == "sum":
        assert np.all(on[1][0][..., 3] == 2 * spp)


def test_random_rays_match_oracle(rt, port, scenes, table):
    rng = np.random.RandomState(11)
    n = 20000
    o = rng.uniform(-6, 6, (n, 3)).astype(np.float32)
    v = rng.normal(size=(n, 3))
    v = (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)
    rays = np.concatenate([o, v], 1)
    for name in ("default", "c3"):
        arr, _ = scenes[name]
        d = op.make_desc(arr)
        want = [port.intersect(d, rays[i, :3], rays[i, 3:]) for i in range(0, n, 10)]
        want_id = np.array([h.id for _, h in want], np.uint32)
        want_t = np.array([h.t for _, h in want], np.float32)
        with make_ctx(rt, scenes, name, 64, 64, table, arith=rt.ARITH_STRICT) as ctx:
            ids, ts = ctx.trace_rays(rays[::10])
            assert np.array_equal(ids, want_id) and np.array_equal(bits(ts), bits(want_t))
            ctx.set_params(arith=rt.ARITH_FAST)
            ids_f, _ = ctx.trace_rays(rays[::10])
            assert (ids_f != want_id).mean() < 1e-3


def test_errors_are_reported_not_fatal(rt, scenes, table):
    arr, tex = scenes["default"]
    from rtb200 import fixtures as fx
    ctx = rt.Context(32, 32)
    with pytest.raises(rt.RtError):
        ctx.render(fx.camera(60, 1.0))               # no scene yet
    bad = dict(arr)
    sph = arr["spheres"].copy()
    sph[0, 20:24] = np.frombuffer(np.uint32(99).tobytes(), np.uint8)   # material id out of range
    bad["spheres"] = sph
    with pytest.raises(rt.RtError) as e:
        ctx.upload_scene(bad)
    assert "material" in str(e.value)
    ctx.upload_scene(arr)
    with pytest.raises(rt.RtError):
        ctx.render(fx.camera(60, 1.0))               # table mode without a table
    ctx.close()


@pytest.mark.parametrize("arith", ["strict", "fast"])
@pytest.mark.parametrize("name,W,H,depth", [("default", 640, 360, 30), ("c3", 480, 270, 30), ("default", 203, 131, 3)])
def test_path_heads_change_nothing(rt, scenes, table, name, W, H, depth, arith):
    """Memoising the sample-independent head of every path (RT_HEAD_ON) must not change one bit of the strict image,
    nor the number of path segments; the executed queries drop.  Fast arm: same statistics (the head is walked in
    Exact arithmetic there, so single pixels may differ from the all-Fused path)."""
    from rtb200 import fixtures as fx
    cam = fx.camera(60, W / H)
    A = rt.ARITH_STRICT if arith == "strict" else rt.ARITH_FAST
    out = {}
    for mode in (rt.HEAD_OFF, rt.HEAD_ON):
        with make_ctx(rt, scenes, name, W, H, table, max_depth=depth, arith=A, rng_mode=rt.RNG_COUNTER, seed=7, path_head=mode) as ctx:
            ctx.render_samples(cam, 0, 6, rt.ACCUM_REFERENCE)
            img = ctx.read_image().copy()
            st = ctx.stats()
            rays, queries = st.rays_last, st.queries_last
            ctx.clear_accum()
            ctx.render_samples(cam, 3, 5, rt.ACCUM_SUM)      # heads cached: no head queries in this call
            acc = ctx.read_accum().copy()
            out[mode] = (img, rays, queries, acc, ctx.stats().rays_last, ctx.stats().queries_last)
    off, on = out[rt.HEAD_OFF], out[rt.HEAD_ON]
    assert off[1] == off[2] and off[4] == off[5]            # without heads every segment is a query
    assert on[2] < on[1] and on[5] < on[4]
    if arith == "strict":
        assert np.array_equal(bits(on[0]), bits(off[0]))
        assert np.array_equal(bits(on[3]), bits(off[3]))
        assert on[1] == off[1] and on[4] == off[4]
    else:
        assert abs(float((on[0][..., :3] ** 2).mean()) - float((off[0][..., :3] ** 2).mean())) < 5e-3
        assert abs(on[1] - off[1]) <= 1e-3 * off[1]
        assert np.all(on[3][..., 3] == 5.0) and np.all(off[3][..., 3] == 5.0)


def test_fused_peer_finalize_equals_reduce_then_finalize(rt, scenes, table):
    """rt_finalize_peers (the fused exchange step of the sample split): three 'ranks' as three contexts on this GPU,
    disjoint global sample ranges, every rank finalizes its slice of rows out of all accumulation buffers into rank
    0's image.  Must equal sum -> rt_finalize of a single context rendering all samples (float summation order aside)."""
    from rtb200 import fixtures as fx
    from rtb200.multigpu import sample_range
    W, H, spp, world = 250, 131, 12, 3
    cam = fx.camera(60, W / H)
    kw = dict(max_depth=30, arith=rt.ARITH_FAST, rng_mode=rt.RNG_COUNTER_TABLE, seed=9)
    ctxs = [make_ctx(rt, scenes, "default", W, H, table, **kw) for _ in range(world)]
    try:
        for r, c in enumerate(ctxs):
            first, n = sample_range(spp, world, r)
            c.clear_accum()
            c.render_samples(cam, first, n, rt.ACCUM_SUM)
        ptrs = [c.accum_device_ptr()[0] for c in ctxs]
        img0 = ctxs[0].image_device_ptr()[0]
        for r, c in enumerate(ctxs):
            c.set_peers_direct(r, ptrs, img0)
        for c in ctxs:
            c.finalize_peers(spp)
        got = ctxs[0].read_image()
        with make_ctx(rt, scenes, "default", W, H, table, **kw) as one:
            one.clear_accum()
            one.render_samples(cam, 0, spp, rt.ACCUM_SUM)
            one.finalize(spp)
            want = one.read_image()
        assert np.all(got[..., 3] == 1.0)
        assert np.abs(got - want).max() <= 2e-6
        with pytest.raises(rt.RtError):
            ctxs[1].close_peers()
            ctxs[1].finalize_peers(spp)
    finally:
        for c in ctxs:
            c.close()


def test_restore_accum_and_image_continue_the_run(rt, scenes, table):
    """rt_restore_image / rt_restore_accum (checkpoint/resume): a second context continues where the first stopped."""
    from rtb200 import fixtures as fx
    W, H = 96, 64
    cam = fx.camera(60, W / H)
    kw = dict(max_depth=30, arith=rt.ARITH_STRICT, rng_mode=rt.RNG_COUNTER_TABLE, seed=2)
    with make_ctx(rt, scenes, "default", W, H, table, **kw) as a, make_ctx(rt, scenes, "default", W, H, table, **kw) as b:
        a.render_samples(cam, 0, 6, rt.ACCUM_REFERENCE)
        want = a.read_image().copy()
        a.clear_accum(); a.render_samples(cam, 0, 6, rt.ACCUM_SUM)
        want_acc = a.read_accum().copy()
        a.render_samples(cam, 0, 2, rt.ACCUM_REFERENCE)
        b.restore_image(a.read_image(), 2)
        for _ in range(4):
            b.render_again(cam)
        assert b.stats().samples_done == 6 and np.array_equal(bits(b.read_image()), bits(want))
        a.clear_accum(); a.render_samples(cam, 0, 2, rt.ACCUM_SUM)
        b.restore_accum(a.read_accum(), 2)
        b.render_samples(cam, 2, 4, rt.ACCUM_SUM)
        assert b.stats().samples_done == 6 and np.array_equal(bits(b.read_accum()), bits(want_acc))
        with pytest.raises(rt.RtError):
            b.restore_image(want, 0)
