"""rtb200.synth — seeded synthetic scenes in the reference's host AoS layouts (include/scene.h:32-81), for the large
BASELINE configs that no scene file of the reference covers:
  C4  closed displaced icosphere mesh (CCW, outward normals) + uniformly placed spheres
  C5  many mixed-material spheres
Everything is generated with numpy RandomState from a documented seed and passed through rt_upload_scene."""
from __future__ import annotations

import numpy as np

MATERIAL_DT = np.dtype({"names": ["type", "color", "extra"], "formats": ["<i4", ("<f4", 4), "<f4"], "offsets": [0, 16, 32], "itemsize": 48})
SPHERE_DT = np.dtype({"names": ["pos", "r", "mat"], "formats": [("<f4", 4), "<f4", "<u4"], "offsets": [0, 16, 20], "itemsize": 32})
PLANE_DT = np.dtype({"names": ["pos", "normal", "mat"], "formats": [("<f4", 4), ("<f4", 4), "<u4"], "offsets": [0, 16, 32], "itemsize": 48})
LENS_DT = np.dtype({"names": ["pos", "p1", "p2", "r1", "r2", "mat"], "formats": [("<f4", 4), ("<f4", 4), ("<f4", 4), "<f4", "<f4", "<u4"],
                    "offsets": [0, 16, 32, 48, 52, 56], "itemsize": 64})
T_REFRACTIVE, T_REFLECTIVE, T_DIELECTRIC, T_DIFFUSE, T_TEXTURED, T_LIGHT = range(6)


def _bytes(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint8).reshape(a.shape[0], -1) if a.shape[0] else np.zeros((0, max(a.dtype.itemsize, 1)), np.uint8)


def materials(specs):
    m = np.zeros(len(specs), MATERIAL_DT)
    for i, (t, col, extra) in enumerate(specs):
        m[i] = (t, (col[0], col[1], col[2], 0.0), extra)
    return m


def icosphere(subdiv):
    """Unit icosphere, counter-clockwise seen from outside: 20 * 4**subdiv faces."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    for _ in range(subdiv):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        es = np.sort(e, axis=1)
        uniq, inv = np.unique(es, axis=0, return_inverse=True)
        mid = v[uniq[:, 0]] + v[uniq[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mid])
        n = len(f)
        ab, bc, ca = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        f = np.concatenate([np.stack([f[:, 0], ab, ca], 1), np.stack([f[:, 1], bc, ab], 1), np.stack([f[:, 2], ca, bc], 1), np.stack([ab, bc, ca], 1)])
    return v, f


def scene(n_spheres=1000, mesh_subdiv=4, seed=1, sphere_box=((-12, -6, 4), (12, 1.2, 40)), radii=(0.05, 0.2), mesh_center=(0.0, -1.0, 9.0),
          mesh_radius=3.0, displacement=0.15, with_plane=True, mesh_material=3, n_sphere_materials=6):
    """-> (arrays dict for rt_upload_scene, description).  Materials: 0 light, 1 reflective, 2 refractive, 3 diffuse grey,
    4 dielectric, 5 diffuse red, 6 diffuse green, 7 floor.  A big light sphere hangs above (world -y is up)."""
    rng = np.random.RandomState(seed)
    mats = materials([(T_LIGHT, (1, 1, 1), 0), (T_REFLECTIVE, (0.95, 0.95, 1), 0.85), (T_REFRACTIVE, (1, 1, 1), 1.3),
                      (T_DIFFUSE, (0.8, 0.8, 0.8), 0.9), (T_DIELECTRIC, (1, 1, 1), 1.5), (T_DIFFUSE, (0.9, 0.25, 0.2), 0.9),
                      (T_DIFFUSE, (0.2, 0.8, 0.3), 0.9), (T_DIFFUSE, (0.75, 0.75, 0.7), 1.0)])
    sph = np.zeros(n_spheres + 1, SPHERE_DT)
    sph[0] = ((0, -400, 15, 0), 300.0, 0)
    lo, hi = np.array(sphere_box[0]), np.array(sphere_box[1])
    pos = rng.uniform(lo, hi, (n_spheres, 3))
    rad = rng.uniform(radii[0], radii[1], n_spheres)
    for i in range(n_spheres):
        sph[i + 1] = ((pos[i, 0], pos[i, 1], pos[i, 2], 0), rad[i], 1 + (i % n_sphere_materials) if n_sphere_materials < 7 else i % 7)
    pla = np.zeros(1 if with_plane else 0, PLANE_DT)
    if with_plane:
        pla[0] = ((0, 1.5, 0, 0), (0, 1, 0, 0), 7)
    arrays = {"materials": _bytes(mats), "spheres": _bytes(sph), "planes": _bytes(pla), "lenses": np.zeros((0, 64), np.uint8)}
    if mesh_subdiv is None:
        arrays.update(vertices=np.zeros((0, 16), np.uint8), uvs=np.zeros((0, 8), np.uint8), indices=np.zeros((0, 4), np.uint8),
                      meshes=np.zeros((0, 16), np.uint8), models=np.zeros((0, 12), np.uint8))
        n_tri = 0
    else:
        v, f = icosphere(mesh_subdiv)
        # smooth radial displacement keeps the surface closed and star-shaped
        bump = 1.0 + displacement * (np.sin(5 * v[:, 0] + 1.3) * np.sin(4 * v[:, 1] + 0.4) * np.sin(6 * v[:, 2] + 2.1))
        v = v * bump[:, None] * mesh_radius + np.asarray(mesh_center)
        verts = np.zeros((len(v), 4), np.float32)
        verts[:, :3] = v
        uvs = np.stack([0.5 + np.arctan2(v[:, 2] - mesh_center[2], v[:, 0] - mesh_center[0]) / (2 * np.pi),
                        0.5 + (v[:, 1] - mesh_center[1]) / (2.4 * mesh_radius)], 1).astype(np.float32)
        arrays.update(vertices=_bytes(verts), uvs=_bytes(uvs), indices=_bytes(f.astype(np.uint32).reshape(-1, 1)),
                      meshes=_bytes(np.array([[0, 0, len(f), 0xFFFFFFFF]], np.uint32)), models=_bytes(np.array([[0, 1, mesh_material]], np.uint32)))
        n_tri = len(f)
    return arrays, dict(seed=seed, spheres=n_spheres + 1, triangles=n_tri, planes=len(pla))


def c4_scene(subdiv=7, n_spheres=1000, seed=1):
    """BASELINE config C4: ~1 M triangles (20 * 4**7 = 327 680; subdiv 8 = 1 310 720) + 1k spheres."""
    return scene(n_spheres=n_spheres, mesh_subdiv=subdiv, seed=seed)


def c5_scene(n_spheres=100000, seed=2):
    """BASELINE config C5: 100 k mixed-material spheres in a slab, no mesh."""
    return scene(n_spheres=n_spheres, mesh_subdiv=None, seed=seed, sphere_box=((-40, -8, 3), (40, 1.3, 120)), radii=(0.02, 0.1))
