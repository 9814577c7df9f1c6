"""rtb200 — thin ctypes binding of librtb200.so (include/rt_capi.h).

Python is only the test/bench harness language here; the product is the C-ABI library and the C++ classes on
top of it.  There is no fallback: if the shared library is missing or no CUDA device is usable, importing
works (so CPU-only checks can inspect symbols) but creating a Context raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "librtb200.so")

RNG_TABLE, RNG_COUNTER, RNG_COUNTER_TABLE = 0, 1, 2
ARITH_STRICT, ARITH_FAST = 0, 1
ACCUM_REFERENCE, ACCUM_SUM = 0, 1
MESH_FIRST_FRONT, MESH_NEAREST = 0, 1
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_BVH = 0, 1, 2
HEAD_ON, HEAD_OFF = 0, 1
ID_MISS = 0xFFFFFFFF
TABLE_FLOATS = 400000
NOALLOC_MAX_RAYS = 65536


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rt error %d: %s" % (code, msg))
        self.code = code


class SceneDesc(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("materials", "spheres", "planes", "lenses", "vertices", "uvs", "indices", "meshes", "models")] + \
               [(n, C.c_uint32) for n in
                ("n_materials", "n_spheres", "n_planes", "n_lenses", "n_vertices", "n_uvs", "n_indices", "n_meshes",
                 "n_models")]


class Params(C.Structure):
    _fields_ = [("max_depth", C.c_uint32), ("rng_mode", C.c_uint32), ("seed", C.c_uint32), ("arith", C.c_uint32),
                ("mesh_mode", C.c_uint32), ("accel", C.c_uint32), ("path_head", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("rays_total", C.c_uint64), ("rays_last", C.c_uint64), ("queries_last", C.c_uint64), ("ms_last", C.c_double), ("ms_trace_last", C.c_double),
                ("samples_done", C.c_uint32), ("kernel_launches", C.c_uint32), ("n_primitives", C.c_uint32),
                ("bvh_nodes", C.c_uint32)]


_lib = None


def lib():
    """Load librtb200.so (raises if it has not been built: run `make -C opencl-raytracing_b200`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(LIB_PATH + " is missing — build it with __graft_entry__.build(); there is no fallback")
    L = C.CDLL(LIB_PATH)
    vp, u32, ci = C.c_void_p, C.c_uint32, C.c_int
    L.rt_last_error.restype = C.c_char_p
    L.rt_version.restype = C.c_char_p
    sig = {
        "rt_create": [ci, ci, ci, C.POINTER(vp)], "rt_destroy": [vp], "rt_set_stream": [vp, vp],
        "rt_upload_scene": [vp, C.POINTER(SceneDesc)], "rt_upload_textures": [vp, vp, ci, ci, ci],
        "rt_make_random_table": [u32, vp], "rt_set_random_table": [vp, vp, u32],
        "rt_set_region": [vp, ci, ci, ci, ci], "rt_set_pixel_split": [vp, ci, ci], "rt_set_work_counting": [vp, ci], "rt_get_work_counts": [vp, vp], "rt_set_params": [vp, C.POINTER(Params)], "rt_get_params": [vp, C.POINTER(Params)],
        "rt_render": [vp, vp], "rt_render_again": [vp, vp], "rt_render_samples": [vp, vp, u32, u32, u32],
        "rt_clear_accum": [vp], "rt_invalidate_caches": [vp], "rt_finalize": [vp, u32], "rt_read_image": [vp, vp], "rt_read_accum": [vp, vp],
        "rt_read_primary_ids": [vp, vp, vp, vp], "rt_restore_image": [vp, vp, u32], "rt_restore_accum": [vp, vp, u32], "rt_get_stats": [vp, C.POINTER(Stats)],
        "rt_accum_device_ptr": [vp, C.POINTER(vp), C.POINTER(C.c_size_t)],
        "rt_image_device_ptr": [vp, C.POINTER(vp), C.POINTER(C.c_size_t)],
        "rt_set_accum_device_buffer": [vp, vp, C.c_size_t], "rt_set_finalize_target": [vp, vp, C.c_size_t],
        "rt_trace_rays": [vp, vp, u32, vp, vp], "rt_trace_rays_noalloc": [vp, vp, u32, vp, vp], "rt_bounce_rays": [vp, vp, u32, vp, vp, vp, vp, vp, vp],
        "rt_ipc_export": [vp, vp, vp], "rt_ipc_open_peers": [vp, ci, ci, vp, vp],
        "rt_set_peers_direct": [vp, ci, ci, C.POINTER(vp), vp], "rt_finalize_peers": [vp, u32], "rt_finalize_peers_async": [vp, u32], "rt_finalize_peers_barrier_async": [vp, u32], "rt_peer_barrier_status": [vp, C.POINTER(ci)], "rt_close_peers": [vp],
        "rt_measure_fp32_peak": [ci, C.POINTER(C.c_double), C.POINTER(C.c_double)],
        "rt_scene_load": [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(vp)], "rt_scene_get_desc": [vp, C.POINTER(SceneDesc)],
        "rt_scene_get_textures": [vp, C.POINTER(vp), C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)],
        "rt_scene_upload": [vp, vp], "rt_scene_free": [vp],
        "rt_camera_data": [ci, C.c_float, vp, C.c_float, C.c_float, vp], "rt_write_tga": [C.c_char_p, vp, ci, ci],
        "rt_camera_run": [ci, C.c_float, vp, C.c_float, C.c_float, vp, ci, vp],
    }
    for name, args in sig.items():
        f = getattr(L, name)
        f.argtypes = args
        f.restype = ci
    _lib = L
    return L


def _check(code):
    if code != 0:
        raise RtError(code, lib().rt_last_error().decode(errors="replace"))


def make_desc(arr):
    """dict of numpy arrays in the reference's host AoS layouts -> SceneDesc (keeps the arrays alive)."""
    d = SceneDesc()
    keep = {}
    for name in ("materials", "spheres", "planes", "lenses", "vertices", "uvs", "indices", "meshes", "models"):
        a = np.ascontiguousarray(arr[name])
        keep[name] = a
        setattr(d, name, a.ctypes.data if a.size else None)
        setattr(d, "n_" + name, a.shape[0])
    d._keep = keep
    return d


def make_random_table(seed):
    t = np.empty(TABLE_FLOATS, dtype=np.float32)
    _check(lib().rt_make_random_table(seed, t.ctypes.data))
    return t


def measure_fp32_peak(device=0):
    g, mhz = C.c_double(0), C.c_double(0)
    _check(lib().rt_measure_fp32_peak(device, C.byref(g), C.byref(mhz)))
    return g.value, mhz.value


ITEM_BYTES = dict(materials=48, spheres=32, planes=48, lenses=64, vertices=16, uvs=8, indices=4, meshes=16, models=12)


def load_scene_file(scene_path, asset_root="", texture_dir=None):
    """SceneCreator::loadScene + loadTextures through the C++ host classes (no GPU needed).
    -> (arrays dict of (n, itemsize) uint8 copies in the reference's host layouts, float textures [L,H,W,4] or None)."""
    L = lib()
    h = C.c_void_p()
    _check(L.rt_scene_load(scene_path.encode(), (asset_root or "").encode(), texture_dir.encode() if texture_dir else None, C.byref(h)))
    try:
        d = SceneDesc()
        _check(L.rt_scene_get_desc(h, C.byref(d)))
        arrays = {}
        for name, nb in ITEM_BYTES.items():
            n = getattr(d, "n_" + name)
            ptr = getattr(d, name)
            arrays[name] = (np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n * nb,)).reshape(n, nb).copy()
                            if n else np.zeros((0, nb), np.uint8))
        p, w, hh, layers = C.c_void_p(), ci_(), ci_(), ci_()
        _check(L.rt_scene_get_textures(h, C.byref(p), C.byref(w), C.byref(hh), C.byref(layers)))
        tex = None
        if layers.value > 0:
            n = layers.value * hh.value * w.value * 4
            tex = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n,)).reshape(layers.value, hh.value, w.value, 4).copy()
        return arrays, tex
    finally:
        L.rt_scene_free(h)


def ci_():
    return C.c_int(0)


def camera_data(fov, aspect, pos=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0):
    """Camera(fov, aspect, pos, yaw, pitch).transferData() of the C++ host class."""
    out = np.zeros(12, np.float32)
    p = np.asarray(pos, np.float32)
    _check(lib().rt_camera_data(int(fov), float(aspect), p.ctypes.data, float(yaw), float(pitch), out.ctypes.data))
    return out


CAM_FORWARD, CAM_BACK, CAM_LEFT, CAM_RIGHT, CAM_ROTATE, CAM_ZOOM, CAM_FAST, CAM_SLOW, CAM_SIZE = range(9)


def camera_run(fov, aspect, pos, yaw, pitch, ops):
    """Camera(...) followed by the controls in ops = [(code, a, b), ...] (rt_camera_run); returns transferData()."""
    out = np.zeros(12, np.float32)
    p = np.asarray(pos, np.float32)
    o = np.ascontiguousarray(np.asarray(ops, np.float32).reshape(-1, 3))
    _check(lib().rt_camera_run(int(fov), float(aspect), p.ctypes.data, float(yaw), float(pitch), o.ctypes.data, len(o), out.ctypes.data))
    return out


class Context:
    """One renderer on one CUDA device: RayTracer + KernelGL + SceneCreator's device half."""

    def __init__(self, width, height, device=0):
        self.L = lib()
        self.W, self.H = width, height
        h = C.c_void_p()
        _check(self.L.rt_create(width, height, device, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.rt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_stream(self, cuda_stream_ptr):
        _check(self.L.rt_set_stream(self.h, cuda_stream_ptr))

    def upload_scene(self, arrays):
        d = make_desc(arrays)
        _check(self.L.rt_upload_scene(self.h, C.byref(d)))

    def upload_textures(self, tex_f32):
        t = np.ascontiguousarray(tex_f32, dtype=np.float32)
        assert t.ndim == 4 and t.shape[3] == 4
        _check(self.L.rt_upload_textures(self.h, t.ctypes.data, t.shape[2], t.shape[1], t.shape[0]))

    def set_random_table(self, table):
        t = np.ascontiguousarray(table, dtype=np.float32)
        _check(self.L.rt_set_random_table(self.h, t.ctypes.data, t.size))

    def set_params(self, max_depth=30, rng_mode=RNG_TABLE, seed=0, arith=ARITH_STRICT, mesh_mode=MESH_FIRST_FRONT,
                   accel=ACCEL_AUTO, path_head=HEAD_ON):
        p = Params(max_depth, rng_mode, seed, arith, mesh_mode, accel, path_head)
        _check(self.L.rt_set_params(self.h, C.byref(p)))

    def set_region(self, x0=0, y0=0, w=0, h=0):
        """Render only pixels [x0, x0+w) x [y0, y0+h); w <= 0 restores the whole frame."""
        _check(self.L.rt_set_region(self.h, x0, y0, w, h))

    def set_pixel_split(self, n_parts=1, part=0):
        """Render only the pixel groups g with g % n_parts == part (128-pixel groups in launch order)."""
        _check(self.L.rt_set_pixel_split(self.h, n_parts, part))

    def set_work_counting(self, on):
        _check(self.L.rt_set_work_counting(self.h, 1 if on else 0))

    def work_counts(self):
        """-> dict of the counted work of the last LBVH render made while counting was on."""
        w = np.zeros(8, np.uint64)
        _check(self.L.rt_get_work_counts(self.h, w.ctypes.data))
        return dict(node_visits=int(w[0]), sphere_tests=int(w[1]), lens_tests=int(w[2]), triangle_tests=int(w[3]),
                    plane_tests=int(w[4]), queries=int(w[5]), flared_node_visits=int(w[6]), flared_queries=int(w[7]))

    @staticmethod
    def _cam(cam):
        c = np.ascontiguousarray(cam, dtype=np.float32)
        assert c.size == 12
        return c

    def render(self, cam):
        c = self._cam(cam)
        _check(self.L.rt_render(self.h, c.ctypes.data))

    def render_again(self, cam):
        c = self._cam(cam)
        _check(self.L.rt_render_again(self.h, c.ctypes.data))

    def render_samples(self, cam, first, n, accum_mode=ACCUM_REFERENCE):
        c = self._cam(cam)
        _check(self.L.rt_render_samples(self.h, c.ctypes.data, first, n, accum_mode))

    def clear_accum(self):
        _check(self.L.rt_clear_accum(self.h))

    def invalidate_caches(self):
        _check(self.L.rt_invalidate_caches(self.h))

    def finalize(self, spp):
        _check(self.L.rt_finalize(self.h, spp))

    def read_image(self, out=None):
        img = out if out is not None else np.empty((self.H, self.W, 4), dtype=np.float32)
        _check(self.L.rt_read_image(self.h, img.ctypes.data))
        return img

    def read_accum(self, out=None):
        img = out if out is not None else np.empty((self.H, self.W, 4), dtype=np.float32)
        _check(self.L.rt_read_accum(self.h, img.ctypes.data))
        return img

    def restore_image(self, img, samples):
        a = np.ascontiguousarray(img, dtype=np.float32)
        assert a.shape == (self.H, self.W, 4)
        _check(self.L.rt_restore_image(self.h, a.ctypes.data, samples))

    def restore_accum(self, acc, samples):
        a = np.ascontiguousarray(acc, dtype=np.float32)
        assert a.shape == (self.H, self.W, 4)
        _check(self.L.rt_restore_accum(self.h, a.ctypes.data, samples))

    def primary_ids(self, cam):
        c = self._cam(cam)
        ids = np.empty((self.H, self.W), dtype=np.uint32)
        ts = np.empty((self.H, self.W), dtype=np.float32)
        _check(self.L.rt_read_primary_ids(self.h, c.ctypes.data, ids.ctypes.data, ts.ctypes.data))
        return ids, ts

    def stats(self):
        s = Stats()
        _check(self.L.rt_get_stats(self.h, C.byref(s)))
        return s

    def accum_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(self.L.rt_accum_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def image_device_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(self.L.rt_image_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def set_accum_device_buffer(self, ptr, nbytes):
        _check(self.L.rt_set_accum_device_buffer(self.h, ptr, nbytes))

    def set_finalize_target(self, ptr, nbytes):
        """rt_finalize / rt_finalize_peers* write the image to this device-accessible address (e.g. mapped pinned host memory)."""
        _check(self.L.rt_set_finalize_target(self.h, ptr, nbytes))

    def ipc_export(self):
        """-> (accum handle, image handle), 64 bytes each (CUDA IPC)."""
        a, i = C.create_string_buffer(64), C.create_string_buffer(64)
        _check(self.L.rt_ipc_export(self.h, a, i))
        return a.raw, i.raw

    def ipc_open_peers(self, rank, accum_handles, image_handle_rank0):
        blob = b"".join(accum_handles)
        _check(self.L.rt_ipc_open_peers(self.h, len(accum_handles), rank, blob, image_handle_rank0))

    def set_peers_direct(self, rank, accum_ptrs, image_ptr_rank0):
        arr = (C.c_void_p * len(accum_ptrs))(*accum_ptrs)
        _check(self.L.rt_set_peers_direct(self.h, len(accum_ptrs), rank, arr, image_ptr_rank0))

    def finalize_peers(self, spp, sync=True):
        _check((self.L.rt_finalize_peers if sync else self.L.rt_finalize_peers_async)(self.h, spp))

    def finalize_peers_barrier(self, spp):
        """Fused reduce + finalize bracketed by the device-side rank barrier; enqueue only."""
        _check(self.L.rt_finalize_peers_barrier_async(self.h, spp))

    def peer_barrier_timed_out(self):
        v = C.c_int(0)
        _check(self.L.rt_peer_barrier_status(self.h, C.byref(v)))
        return bool(v.value)

    def close_peers(self):
        _check(self.L.rt_close_peers(self.h))

    def trace_rays(self, rays6):
        r = np.ascontiguousarray(rays6, dtype=np.float32).reshape(-1, 6)
        ids = np.empty(r.shape[0], dtype=np.uint32)
        ts = np.empty(r.shape[0], dtype=np.float32)
        _check(self.L.rt_trace_rays(self.h, r.ctypes.data, r.shape[0], ids.ctypes.data, ts.ctypes.data))
        return ids, ts

    def trace_rays_noalloc(self, rays6):
        """rt_trace_rays_noalloc: strict arithmetic, static device scratch, at most NOALLOC_MAX_RAYS rays."""
        r = np.ascontiguousarray(rays6, dtype=np.float32).reshape(-1, 6)
        ids = np.empty(r.shape[0], dtype=np.uint32)
        ts = np.empty(r.shape[0], dtype=np.float32)
        _check(self.L.rt_trace_rays_noalloc(self.h, r.ctypes.data, r.shape[0], ids.ctypes.data, ts.ctypes.data))
        return ids, ts

    def bounce_rays(self, rays6, sample, bounce, gx, gy):
        r = np.array(rays6, dtype=np.float32).reshape(-1, 6)
        n = r.shape[0]
        a = [np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.uint32), (n,))) for v in (sample, bounce, gx, gy)]
        col = np.empty((n, 3), dtype=np.float32)
        ended = np.empty(n, dtype=np.uint8)
        _check(self.L.rt_bounce_rays(self.h, r.ctypes.data, n, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data,
                                     a[3].ctypes.data, col.ctypes.data, ended.ctypes.data))
        return r, col, ended.astype(bool)
