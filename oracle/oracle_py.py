"""oracle/oracle_py.py — TEST INFRASTRUCTURE ONLY: ctypes bindings of the checkers.

  RefOracle  : oracle/_ref/libref_{strict,fast}.so — the reference's own kernel text compiled as C++
  PortOracle : oracle/lib/liboracle_port.so        — the plain-C restatement (oracle/rt_oracle.c)

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TABLE_FLOATS = 400000


class SceneDesc(C.Structure):
    """Layout shared by ref_scene_desc (oracle) and rt_scene_desc (include/rt_capi.h)."""
    _fields_ = [(n, C.c_void_p) for n in
                ("materials", "spheres", "planes", "lenses", "vertices", "uvs", "indices", "meshes", "models")] + \
               [(n, C.c_uint32) for n in
                ("n_materials", "n_spheres", "n_planes", "n_lenses", "n_vertices", "n_uvs", "n_indices", "n_meshes",
                 "n_models")]


def make_desc(arr):
    """arr: dict of numpy arrays in the reference AoS layouts (see scene_py.Scene.arrays)."""
    keep = {}
    d = SceneDesc()
    for name in ("materials", "spheres", "planes", "lenses", "vertices", "uvs", "indices", "meshes", "models"):
        a = np.ascontiguousarray(arr[name])
        keep[name] = a
        setattr(d, name, a.ctypes.data if a.size else None)
        setattr(d, "n_" + name, a.shape[0])
    d._keep = keep
    return d


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class RefOracle:
    def __init__(self, flavour="strict"):
        path = os.path.join(HERE, "_ref", "libref_%s.so" % flavour)
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (build with `make -C oracle ref` where /root/reference exists)")
        L = self.lib = C.CDLL(path)
        L.ref_render.restype = C.c_double
        L.ref_render.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                 C.POINTER(C.c_uint64)]
        L.ref_make_table.argtypes = [C.c_uint32, C.c_void_p]
        L.ref_hit.restype = C.c_int
        L.ref_hit.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_primary_ids.restype = C.c_int
        L.ref_primary_ids.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_identify.restype = C.c_uint32
        L.ref_identify.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_bounce.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
        L.ref_build_info.restype = C.c_char_p

    def build_info(self):
        return self.lib.ref_build_info().decode()

    def make_table(self, seed):
        t = np.empty(TABLE_FLOATS, dtype=np.float32)
        self.lib.ref_make_table(seed, t.ctypes.data)
        return t

    def render(self, desc, tex, cam, table, W, H, first, n, depth=30, threads=None, image=None):
        """Returns (image RGBA f32 [H,W,4] = sqrt(running mean), rays, seconds)."""
        tex = _tex(tex)
        if image is None:
            image = np.zeros((H, W, 4), dtype=np.float32)
        rays = C.c_uint64(0)
        cam = np.ascontiguousarray(cam, dtype=np.float32)
        secs = self.lib.ref_render(C.byref(desc), tex.ctypes.data, tex.shape[2], tex.shape[1], tex.shape[0],
                                   cam.ctypes.data, table.ctypes.data, image.ctypes.data, W, H, first, n, depth,
                                   threads or os.cpu_count(), C.byref(rays))
        return image, rays.value, secs

    def hit(self, desc, origin, direction):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        out = np.zeros(9, dtype=np.float32)
        ids = np.zeros(2, dtype=np.uint32)
        h = self.lib.ref_hit(C.byref(desc), o.ctypes.data, d.ctypes.data, out.ctypes.data, ids.ctypes.data)
        return bool(h), out, ids

    def identify(self, desc, origin, direction):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        t = C.c_float(0)
        i = self.lib.ref_identify(C.byref(desc), o.ctypes.data, d.ctypes.data, C.byref(t))
        return i, t.value

    def primary_ids(self, desc, cam, W, H):
        ids = np.zeros((H, W), dtype=np.uint32)
        ts = np.zeros((H, W), dtype=np.float32)
        cam = np.ascontiguousarray(cam, dtype=np.float32)
        bad = self.lib.ref_primary_ids(C.byref(desc), cam.ctypes.data, W, H, ids.ctypes.data, ts.ctypes.data)
        return ids, ts, bad

    def bounce(self, desc, tex, table, ray6, s_seed, gx, gy):
        tex = _tex(tex)
        ray = np.array(ray6, dtype=np.float32)
        col = np.zeros(3, dtype=np.float32)
        self.lib.ref_bounce(C.byref(desc), tex.ctypes.data, tex.shape[2], tex.shape[1], tex.shape[0],
                            table.ctypes.data, ray.ctypes.data, s_seed, gx, gy, col.ctypes.data)
        return ray, col


def _tex(tex):
    if tex is None:
        return np.zeros((1, 1, 1, 4), dtype=np.float32)
    assert tex.dtype == np.float32 and tex.ndim == 4 and tex.shape[3] == 4
    return np.ascontiguousarray(tex)


class OrcParams(C.Structure):
    _fields_ = [("max_depth", C.c_uint32), ("rng_mode", C.c_uint32), ("seed", C.c_uint32),
                ("accum_mode", C.c_uint32), ("mesh_mode", C.c_uint32), ("threads", C.c_int32)]


class OrcHit(C.Structure):
    _fields_ = [("t", C.c_float), ("id", C.c_uint32), ("p", C.c_float * 3), ("n", C.c_float * 3),
                ("uv", C.c_float * 2), ("mat_id", C.c_uint32), ("texture_id", C.c_uint32)]


RNG_TABLE, RNG_COUNTER, RNG_COUNTER_TABLE = 0, 1, 2
ACCUM_REFERENCE, ACCUM_SUM = 0, 1
MESH_FIRST_FRONT, MESH_NEAREST = 0, 1


class PortOracle:
    def __init__(self):
        path = os.path.join(HERE, "lib", "liboracle_port.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (build with `make -C oracle port`)")
        L = self.lib = C.CDLL(path)
        L.orc_render.restype = C.c_double
        L.orc_render.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(OrcParams),
                                 C.POINTER(C.c_uint64)]
        L.orc_render_window.restype = C.c_double
        L.orc_render_window.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                        C.POINTER(OrcParams), C.POINTER(C.c_uint64)]
        L.orc_intersect.restype = C.c_int
        L.orc_intersect.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(OrcHit)]
        L.orc_primary_ids.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_void_p,
                                      C.c_void_p]
        L.orc_bounce.restype = C.c_int
        L.orc_bounce.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                 C.POINTER(OrcParams), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_void_p]
        L.orc_rng_mix32.restype = C.c_uint32
        L.orc_rng_mix32.argtypes = [C.c_uint32]
        L.orc_rng_bounce_key.restype = C.c_uint32
        L.orc_rng_bounce_key.argtypes = [C.c_uint32] * 4
        L.orc_rng_scatter_vec.argtypes = [C.c_uint32, C.c_void_p]
        L.orc_rng_uniform.restype = C.c_float
        L.orc_rng_uniform.argtypes = [C.c_uint32]
        L.orc_make_table.argtypes = [C.c_uint32, C.c_void_p]

    @staticmethod
    def params(max_depth=30, rng_mode=RNG_TABLE, seed=0, accum_mode=ACCUM_REFERENCE, mesh_mode=MESH_FIRST_FRONT,
               threads=None):
        return OrcParams(max_depth, rng_mode, seed, accum_mode, mesh_mode, threads or os.cpu_count())

    def make_table(self, seed):
        t = np.empty(TABLE_FLOATS, dtype=np.float32)
        self.lib.orc_make_table(seed, t.ctypes.data)
        return t

    def render(self, desc, tex, cam, table, W, H, first, n, prm=None, image=None):
        tex = _tex(tex)
        prm = prm or self.params()
        if image is None:
            image = np.zeros((H, W, 4), dtype=np.float32)
        rays = C.c_uint64(0)
        cam = np.ascontiguousarray(cam, dtype=np.float32)
        tp = table.ctypes.data if table is not None else None
        secs = self.lib.orc_render(C.byref(desc), tex.ctypes.data, tex.shape[2], tex.shape[1], tex.shape[0],
                                   cam.ctypes.data, tp, image.ctypes.data, W, H, first, n, C.byref(prm),
                                   C.byref(rays))
        return image, rays.value, secs

    def render_window(self, desc, tex, cam, table, W, H, x0, y0, w, h, first, n, prm=None):
        """Pixels [x0, x0+w) x [y0, y0+h) of the W x H frame -> (image [h, w, 4], rays, seconds)."""
        tex = _tex(tex)
        prm = prm or self.params()
        image = np.zeros((h, w, 4), dtype=np.float32)
        rays = C.c_uint64(0)
        cam = np.ascontiguousarray(cam, dtype=np.float32)
        tp = table.ctypes.data if table is not None else None
        secs = self.lib.orc_render_window(C.byref(desc), tex.ctypes.data, tex.shape[2], tex.shape[1], tex.shape[0],
                                          cam.ctypes.data, tp, image.ctypes.data, W, H, x0, y0, w, h, first, n,
                                          C.byref(prm), C.byref(rays))
        return image, rays.value, secs

    def intersect(self, desc, origin, direction, mesh_mode=MESH_FIRST_FRONT):
        o = np.ascontiguousarray(origin, dtype=np.float32)
        d = np.ascontiguousarray(direction, dtype=np.float32)
        h = OrcHit()
        hit = self.lib.orc_intersect(C.byref(desc), o.ctypes.data, d.ctypes.data, mesh_mode, C.byref(h))
        return bool(hit), h

    def primary_ids(self, desc, cam, W, H, mesh_mode=MESH_FIRST_FRONT):
        ids = np.zeros((H, W), dtype=np.uint32)
        ts = np.zeros((H, W), dtype=np.float32)
        cam = np.ascontiguousarray(cam, dtype=np.float32)
        self.lib.orc_primary_ids(C.byref(desc), cam.ctypes.data, W, H, mesh_mode, ids.ctypes.data, ts.ctypes.data)
        return ids, ts

    def bounce(self, desc, tex, table, prm, ray6, sample, bounce, gx, gy, W):
        tex = _tex(tex)
        ray = np.array(ray6, dtype=np.float32)
        col = np.zeros(3, dtype=np.float32)
        tp = table.ctypes.data if table is not None else None
        ended = self.lib.orc_bounce(C.byref(desc), tex.ctypes.data, tex.shape[2], tex.shape[1], tex.shape[0], tp,
                                    C.byref(prm), ray.ctypes.data, sample, bounce, gx, gy, W, col.ctypes.data)
        return ray, col, bool(ended)

    def scatter_vec(self, hb):
        v = np.zeros(3, dtype=np.float32)
        self.lib.orc_rng_scatter_vec(hb, v.ctypes.data)
        return v
