"""Known-answer tests of the oracle's primitive routines from analytic geometry, exercising the reference's
conventions: range [0.001, 1000], near-then-far sphere root, double-sided plane normal,
triangle epsilon and barycentric bounds, CCW normal, first-front-facing mesh rule, strict-< tie breaks."""
import numpy as np
import pytest

import oracle_py as op
import scene_py as sp


def scene(build):
    s = sp.Scene()
    s.add_material(sp.T_DIFFUSE, (1, 1, 1), 1.0)
    s.add_material(sp.T_LIGHT, (1, 1, 1), 0.0)
    build(s)
    return op.make_desc(s.arrays())


def hit(port, d, o, v, mode=op.MESH_FIRST_FRONT):
    return port.intersect(d, np.array(o, np.float32), np.array(v, np.float32), mode)


def test_sphere_near_and_far_root(port):
    d = scene(lambda s: s.add_sphere((0, 0, 5), 1.0, 0))
    ok, h = hit(port, d, (0, 0, 0), (0, 0, 1))
    assert ok and h.id == (1 << 28) and h.t == pytest.approx(4.0) and tuple(h.n) == pytest.approx((0, 0, -1))
    ok, h = hit(port, d, (0, 0, 5), (0, 0, 1))          # from the centre: near root -1 rejected, far root 1
    assert ok and h.t == pytest.approx(1.0) and tuple(h.n) == pytest.approx((0, 0, 1))   # outward normal
    ok, _ = hit(port, d, (0, 0, 0), (0, 0, -1))
    assert not ok
    ok, _ = hit(port, d, (0, 1.0, 0), (0, 0, 1))         # tangent: discriminant == 0 is a miss (dis > 0)
    assert not ok


def test_range_limits(port):
    d = scene(lambda s: s.add_plane((0, 0, 0), (0, 0, 1), 0))
    assert hit(port, d, (0, 0, -0.002), (0, 0, 1))[0]
    assert not hit(port, d, (0, 0, -0.0005), (0, 0, 1))[0]          # t < 0.001
    assert not hit(port, d, (0, 0, -1000.0), (0, 0, 1))[0]          # t == 1000 is in range but never < 1000
    assert hit(port, d, (0, 0, -999.0), (0, 0, 1))[0]


def test_plane_normal_faces_the_ray(port):
    d = scene(lambda s: s.add_plane((0, 5, 0), (0, 1, 0), 0))
    ok, h = hit(port, d, (0, 0, 0), (0, 1, 0))
    assert ok and h.t == pytest.approx(5.0) and tuple(h.n) == pytest.approx((0, -1, 0))
    ok, h = hit(port, d, (0, 10, 0), (0, -1, 0))
    assert ok and tuple(h.n) == pytest.approx((0, 1, 0))
    assert not hit(port, d, (0, 0, 0), (1, 0, 0))[0]                # parallel: b/0 = inf


def test_ties_first_in_list_wins(port):
    def build(s):
        s.add_sphere((0, 0, 5), 1.0, 0)
        s.add_sphere((0, 0, 5), 1.0, 1)
        s.add_plane((0, 0, 4), (0, 0, 1), 1)
    d = scene(build)
    ok, h = hit(port, d, (0, 0, 0), (0, 0, 1))
    assert ok and h.id == (1 << 28) | 0 and h.mat_id == 0            # equal t: sphere 0 beats sphere 1 and the plane


def _tri_scene(verts, faces, mat=0):
    def build(s):
        s.vertices += [np.array(v, np.float32) for v in verts]
        s.uvs += [np.array([v[0], v[1]], np.float32) for v in verts]
        s.indices += [i for f in faces for i in f]
        s.meshes.append((0, 0, len(faces), 0xFFFFFFFF))
        s.models.append((0, 1, mat))
    return scene(build)


def test_triangle_ccw_front_face_only(port):
    d = _tri_scene([(0, 0, 5), (1, 0, 5), (0, 1, 5)], [(0, 2, 1)])       # normal (0,0,-1): faces the origin
    ok, h = hit(port, d, (0.25, 0.25, 0), (0, 0, 1))
    assert ok and h.id == (4 << 28) and h.t == pytest.approx(5.0) and tuple(h.n) == pytest.approx((0, 0, -1))
    assert tuple(h.uv) == pytest.approx((0.25, 0.25))
    assert not hit(port, d, (0.25, 0.25, 10), (0, 0, -1))[0]              # back face is never hit
    assert not hit(port, d, (0.75, 0.75, 0), (0, 0, 1))[0]                # u + v > 1
    assert hit(port, d, (0.0, 0.0, 0), (0, 0, 1))[0]                      # corner: u = v = 0 accepted


def test_mesh_first_front_facing_vs_nearest(port):
    # two front-facing triangles along +z; the FAR one comes first in face order
    verts = [(-1, -1, 8), (1, -1, 8), (0, 1, 8), (-1, -1, 4), (1, -1, 4), (0, 1, 4)]
    d = _tri_scene(verts, [(0, 2, 1), (3, 5, 4)])
    ok, h = hit(port, d, (0, 0, 0), (0, 0, 1), op.MESH_FIRST_FRONT)
    assert ok and h.id == (4 << 28) | 0 and h.t == pytest.approx(8.0)     # reference rule: first in order
    ok, h = hit(port, d, (0, 0, 0), (0, 0, 1), op.MESH_NEAREST)
    assert ok and h.id == (4 << 28) | 1 and h.t == pytest.approx(4.0)     # BVH-compatible rule: nearest


def test_lens_cases(port):
    d = scene(lambda s: s.add_lens((0, 0, 5), (0, 0, 1), 2.0, 2.0, 1.0, 0))
    k = np.sqrt(3.0)                                    # p1 = (0,0,5+k), p2 = (0,0,5-k); lens spans z in [3+k, 7-k]
    ok, h = hit(port, d, (0, 0, 0), (0, 0, 1))
    assert ok and h.id == (3 << 28) and h.t == pytest.approx(3 + k, rel=1e-6)
    assert tuple(h.n) == pytest.approx((0, 0, -1), abs=1e-6)
    ok, h = hit(port, d, (0, 0, 5), (0, 0, 1))           # from inside: exits through the nearer far root
    assert ok and h.t == pytest.approx(2 - k, rel=1e-5) and tuple(h.n) == pytest.approx((0, 0, 1), abs=1e-6)
    assert not hit(port, d, (0, 1.5, 0), (0, 0, 1))[0]   # passes both spheres but not their intersection
    assert not hit(port, d, (0, 0, 10), (0, 0, 1))[0]    # behind the ray


def test_counter_rng_known_answers(port, table):
    """Spec vectors of the counter generator and its sampling law: it must match the law of
    the reference's misaligned table read (kernels/raytracer.cl:109-117): E|v|^2 = 3/5 and P(|v| > 1) ~ 8.4 %
    (a true in-ball sampler would have P = 0)."""
    L = port.lib
    assert L.orc_rng_mix32(0) == 0 and L.orc_rng_mix32(1) == 0x688990C0 and L.orc_rng_mix32(0xDEADBEEF) == 0xE628C683
    assert L.orc_rng_bounce_key(9, 5, 3, 2) == 0x1BD6B6D1
    assert np.allclose(port.scatter_vec(0x12345678), [-0.07291686534881592, -0.4501066207885742, -0.5408046245574951], rtol=0, atol=1e-7)
    assert L.orc_rng_uniform(0x12345678) == pytest.approx(0.9605576992034912, abs=1e-9)
    n = 40000
    v = np.array([port.scatter_vec(L.orc_rng_bounce_key(9, i, 0, 0)) for i in range(n)])
    idx = np.random.RandomState(0).randint(0, 100000, n)
    tv = np.stack([table[idx], table[idx + 1], table[idx + 2]], 1)
    r2, r2t = (v * v).sum(1), (tv * tv).sum(1)
    assert abs(r2.mean() - 0.6) < 0.01 and abs(r2t.mean() - 0.6) < 0.01
    assert abs((r2 > 1).mean() - (r2t > 1).mean()) < 0.01 and (r2 > 1).mean() > 0.06
    assert np.abs(v.mean(0)).max() < 0.02 and np.abs(np.abs(v).mean(0) - np.abs(tv).mean(0)).max() < 0.01
    u = np.array([L.orc_rng_uniform(L.orc_rng_bounce_key(9, i, 1, 2)) for i in range(n)])
    assert 0.0 <= u.min() and u.max() < 1.0 and abs(u.mean() - 0.5) < 0.01
