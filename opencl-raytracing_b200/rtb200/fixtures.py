"""rtb200.fixtures — load a packed scene (rtb200/data/scene_*.npz: package data, the parsed default scene of the reference
and the C3 cube scene in the reference's host layouts) into the arrays rt_upload_scene takes, and prepare the inputs a host application would prepare: float textures (stb_image's LDR->float law,
as src/scene.cpp:158 obtains them), the random table and the Camera::transferData() vector."""
from __future__ import annotations

import math
import os

import numpy as np

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
ARRAY_NAMES = ("materials", "spheres", "planes", "lenses", "vertices", "uvs", "indices", "meshes", "models")
ITEM_BYTES = dict(materials=48, spheres=32, planes=48, lenses=64, vertices=16, uvs=8, indices=4, meshes=16, models=12)


def load_scene_npz(path):
    """-> (arrays dict of (n, itemsize) uint8 views in the reference's host layouts, RGBA8 textures or None)."""
    z = np.load(path)
    arrays = {}
    for k in ARRAY_NAMES:
        a = np.ascontiguousarray(z[k])
        if a.size == 0:
            a = np.zeros((0, ITEM_BYTES[k]), np.uint8)
        assert a.dtype == np.uint8 and a.shape[1] == ITEM_BYTES[k], (k, a.shape)
        arrays[k] = a
    tex8 = z["textures_rgba8"] if "textures_rgba8" in z.files else None
    return arrays, tex8


def textures_to_float(rgba8):
    """8-bit RGBA -> float RGBA the way stbi_loadf does: pow(c/255, 2.2) on RGB, a/255 on alpha."""
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.powf.restype = ctypes.c_float
    libm.powf.argtypes = [ctypes.c_float, ctypes.c_float]
    c = (np.arange(256, dtype=np.float32) / np.float32(255.0)).astype(np.float32)
    rgb = np.array([libm.powf(float(x), 2.2) for x in c], dtype=np.float32)
    out = np.empty(rgba8.shape, dtype=np.float32)
    out[..., :3] = rgb[rgba8[..., :3]]
    out[..., 3] = c[rgba8[..., 3]]
    return out


def default_scene():
    return load_scene_npz(os.path.join(DATA, "scene_default.npz"))


def c3_scene():
    return load_scene_npz(os.path.join(DATA, "scene_c3.npz"))


def camera(fov_deg, aspect, pos=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0):
    """Camera(fov, aspect, pos, yaw, pitch)->transferData() (src/camera.cpp:26-51,94-118), float32 steps."""
    f = np.float32
    fov = f(int(fov_deg))
    angle = f(np.float64(fov) * math.pi / 180.0)
    hh = f(math.tan(float(angle * f(0.5))))
    hw = f(aspect) * hh
    rad = f(0.01745329251994329576923690768489)
    rp, ry = f(pitch) * rad, f(yaw) * rad
    w = np.array([math.cos(rp) * math.sin(ry), math.sin(rp), math.cos(rp) * math.cos(ry)], dtype=f)
    w = (w * (f(1.0) / np.sqrt((w * w).sum(dtype=f)))).astype(f)
    up = np.array([0.0, -1.0, 0.0], dtype=f)
    u = np.cross(w, up).astype(f)
    u = (u * (f(1.0) / np.sqrt((u * u).sum(dtype=f)))).astype(f)
    v = np.cross(u, w).astype(f)
    llc = w - (hw * u + hh * v)
    return np.concatenate([np.asarray(pos, dtype=f), llc, f(2.0) * hw * u, f(2.0) * hh * v]).astype(f)
