"""oracle/scene_py.py — TEST INFRASTRUCTURE ONLY (never imported by the product).

Python/numpy restatement of the reference's *host-side* scene preparation, used to build the inputs
of the oracle and to cross-check the product's C++ SceneCreator.  Citations are into /root/reference.

  parse_scene_file   src/scene.cpp:297-461   .scene grammar, clang left-to-right argument order
  mat4 helpers       GLM translate/rotate/scale/radians as called at src/scene.cpp:348-353
                     (GLM is not vendored in the reference; restated from GLM's published formulas,
                     float32 arithmetic, libm cosf/sinf through ctypes so results match C++ bit for bit)
  load_obj           Assimp ReadFile(aiProcess_Triangulate|aiProcess_FlipUVs), src/scene.cpp:192-295:
                     one vertex per face corner, fan triangulation (0,1,2),(0,2,3), v -> 1-v
                     (Assimp is not vendored: documented behaviour, "parity unpinned" at this boundary)
  load_texture       stbi_loadf, src/scene.cpp:158: 8-bit -> pow(c/255, 2.2) for RGB, a/255 for alpha
                     (stb_image not vendored: restated from its documented defaults)
  add_lens           src/scene.cpp:122-143
  camera_data        src/camera.cpp:26-51,94-118
"""
from __future__ import annotations

import ctypes
import ctypes.util
import math
import os
import re

import numpy as np

f32 = np.float32
MISS = 0xFFFFFFFF

# enum MatType (include/scene.h:30)
T_REFRACTIVE, T_REFLECTIVE, T_DIELECTRIC, T_DIFFUSE, T_TEXTURED, T_LIGHT = range(6)
MAT_NAMES = {"reflective": T_REFLECTIVE, "refractive": T_REFRACTIVE, "diffuse": T_DIFFUSE,
             "dielectric": T_DIELECTRIC, "light": T_LIGHT, "textured": T_TEXTURED}

# Host AoS layouts (include/scene.h:32-81; cl_float3 is 16 bytes)
MATERIAL_DT = np.dtype({"names": ["type", "color", "extra"], "formats": ["<i4", ("<f4", 4), "<f4"],
                        "offsets": [0, 16, 32], "itemsize": 48})
SPHERE_DT = np.dtype({"names": ["pos", "r", "mat"], "formats": [("<f4", 4), "<f4", "<u4"],
                      "offsets": [0, 16, 20], "itemsize": 32})
PLANE_DT = np.dtype({"names": ["pos", "normal", "mat"], "formats": [("<f4", 4), ("<f4", 4), "<u4"],
                     "offsets": [0, 16, 32], "itemsize": 48})
LENS_DT = np.dtype({"names": ["pos", "p1", "p2", "r1", "r2", "mat"],
                    "formats": [("<f4", 4), ("<f4", 4), ("<f4", 4), "<f4", "<f4", "<u4"],
                    "offsets": [0, 16, 32, 48, 52, 56], "itemsize": 64})
MESH_DT = np.dtype([("vertex_anchor", "<u4"), ("index_anchor", "<u4"), ("face_count", "<u4"), ("texture_ID", "<u4")])
MODEL_DT = np.dtype([("mesh_anchor", "<u4"), ("mesh_count", "<u4"), ("mat", "<u4")])

_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
for _n in ("cosf", "sinf", "tanf", "sqrtf"):
    getattr(_libm, _n).restype = ctypes.c_float
    getattr(_libm, _n).argtypes = [ctypes.c_float]
_libm.powf.restype = ctypes.c_float
_libm.powf.argtypes = [ctypes.c_float, ctypes.c_float]


def _cosf(x): return f32(_libm.cosf(float(x)))
def _sinf(x): return f32(_libm.sinf(float(x)))
def _tanf(x): return f32(_libm.tanf(float(x)))
def _sqrtf(x): return f32(_libm.sqrtf(float(x)))


class SceneError(Exception):
    """The reference prints the message and calls exit(-1) (src/scene.cpp:29-32)."""


# ----------------------------------------------------------------------------- GLM restatement
def mat4_identity():
    return np.eye(4, dtype=f32)  # m[col][row], column-major like glm::mat4


def mat4_translate(m, v):
    r = m.copy()
    r[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3]
    return r


def mat4_scale(m, v):
    r = m.copy()
    r[0] = m[0] * v[0]
    r[1] = m[1] * v[1]
    r[2] = m[2] * v[2]
    return r


def glm_radians(deg):
    return f32(deg) * f32(0.01745329251994329576923690768489)


def mat4_rotate(m, angle, v):
    a = f32(angle)
    c, s = _cosf(a), _sinf(a)
    v = np.asarray(v, dtype=f32)
    inv = f32(1.0) / _sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])  # glm::normalize = v * inversesqrt(dot)
    axis = v * inv
    temp = (f32(1.0) - c) * axis
    rot = np.zeros((3, 3), dtype=f32)
    rot[0][0] = c + temp[0] * axis[0]
    rot[0][1] = temp[0] * axis[1] + s * axis[2]
    rot[0][2] = temp[0] * axis[2] - s * axis[1]
    rot[1][0] = temp[1] * axis[0] - s * axis[2]
    rot[1][1] = c + temp[1] * axis[1]
    rot[1][2] = temp[1] * axis[2] + s * axis[0]
    rot[2][0] = temp[2] * axis[0] + s * axis[1]
    rot[2][1] = temp[2] * axis[1] - s * axis[0]
    rot[2][2] = c + temp[2] * axis[2]
    r = m.copy()
    r[0] = m[0] * rot[0][0] + m[1] * rot[0][1] + m[2] * rot[0][2]
    r[1] = m[0] * rot[1][0] + m[1] * rot[1][1] + m[2] * rot[1][2]
    r[2] = m[0] * rot[2][0] + m[1] * rot[2][1] + m[2] * rot[2][2]
    r[3] = m[3]
    return r


def transform_vertex(v, t):
    """src/scene.cpp:226-232"""
    x, y, z = f32(v[0]), f32(v[1]), f32(v[2])
    return np.array([t[0][0] * x + t[1][0] * y + t[2][0] * z + t[3][0],
                     t[0][1] * x + t[1][1] * y + t[2][1] * z + t[3][1],
                     t[0][2] * x + t[1][2] * y + t[2][2] * z + t[3][2]], dtype=f32)


# ----------------------------------------------------------------------------- OBJ / MTL / PNG
def load_obj(path):
    """Minimal OBJ reader reproducing what the reference consumes from Assimp: a list of meshes
    (one per object/material run), each with per-corner unrolled positions, uvs (v flipped) or None,
    fan-triangulated indices and the diffuse texture path of its material (or None)."""
    positions, uvs = [], []
    meshes = []
    cur = None
    mtl = {}
    cur_mtl = None

    def start_mesh():
        nonlocal cur
        cur = {"pos": [], "uv": [], "idx": [], "has_uv": True, "mtl": cur_mtl}
        meshes.append(cur)

    base = os.path.dirname(path)
    with open(path, "r") as fh:
        for raw in fh:
            line = raw.split("#", 1)[0].strip()
            if not line:
                continue
            tok = line.split()
            if tok[0] == "v":
                positions.append([float(tok[1]), float(tok[2]), float(tok[3])])
            elif tok[0] == "vt":
                uvs.append([float(tok[1]), float(tok[2])])
            elif tok[0] == "mtllib":
                mtl.update(load_mtl(os.path.join(base, line.split(None, 1)[1])))
            elif tok[0] == "usemtl":
                cur_mtl = tok[1]
                if cur is None or cur["idx"]:
                    start_mesh()
                else:
                    cur["mtl"] = cur_mtl
            elif tok[0] in ("o", "g"):
                if cur is None or cur["idx"]:
                    start_mesh()
            elif tok[0] == "f":
                if cur is None:
                    start_mesh()
                corners = []
                for c in tok[1:]:
                    parts = c.split("/")
                    vi = int(parts[0])
                    vi = vi - 1 if vi > 0 else len(positions) + vi
                    ti = None
                    if len(parts) > 1 and parts[1]:
                        ti = int(parts[1])
                        ti = ti - 1 if ti > 0 else len(uvs) + ti
                    base_idx = len(cur["pos"])
                    cur["pos"].append(positions[vi])
                    if ti is None:
                        cur["has_uv"] = False
                        cur["uv"].append([0.0, 0.0])
                    else:
                        cur["uv"].append([f32(uvs[ti][0]), f32(1.0) - f32(uvs[ti][1])])  # aiProcess_FlipUVs, float arithmetic
                    corners.append(base_idx)
                for k in range(1, len(corners) - 1):  # aiProcess_Triangulate on a convex polygon
                    cur["idx"] += [corners[0], corners[k], corners[k + 1]]
    out = []
    for m in meshes:
        if not m["idx"]:
            continue
        out.append({"pos": np.asarray(m["pos"], dtype=f32), "uv": np.asarray(m["uv"], dtype=f32) if m["has_uv"] else None,
                    "idx": np.asarray(m["idx"], dtype=np.uint32), "texture": mtl.get(m["mtl"], {}).get("map_Kd")})
    return out


def load_mtl(path):
    mats, cur = {}, None
    if not os.path.exists(path):
        return mats
    with open(path, "r") as fh:
        for raw in fh:
            line = raw.strip()
            if line.startswith("newmtl"):
                cur = line.split(None, 1)[1]
                mats[cur] = {}
            elif line.startswith("map_Kd") and cur is not None:
                mats[cur]["map_Kd"] = line.split(None, 1)[1]
    return mats


_STB_LUT = None


def stb_ldr_to_hdr_lut():
    """stb_image stbi__ldr_to_hdr: (float)(pow(c / 255.0f, 2.2f) * 1.0f); alpha: c / 255.0f.  stb_image.h is compiled
    inside a C++ translation unit in the reference (src/scene.cpp:16-17), where pow(float, float) is powf."""
    global _STB_LUT
    if _STB_LUT is None:
        c = (np.arange(256, dtype=f32) / f32(255.0)).astype(f32)
        rgb = np.array([_libm.powf(float(x), 2.2) for x in c], dtype=f32)
        _STB_LUT = (rgb, c)
    return _STB_LUT


def texture_from_rgba8(rgba8):
    rgb_lut, a_lut = stb_ldr_to_hdr_lut()
    out = np.empty(rgba8.shape, dtype=f32)
    out[..., :3] = rgb_lut[rgba8[..., :3]]
    out[..., 3] = a_lut[rgba8[..., 3]]
    return out


def load_texture(path):
    from PIL import Image
    im = Image.open(path)
    if im.mode != "RGBA":
        raise SceneError("ERROR: STBimage: TEXTURE HAS A WRONG FORMAT: %d INSTEAD OF 4 (RGBA)" % len(im.getbands()))
    return np.asarray(im, dtype=np.uint8)  # row 0 = top row, no flip


# ----------------------------------------------------------------------------- scene container
class Scene:
    def __init__(self):
        self.materials, self.spheres, self.planes, self.lenses = [], [], [], []
        self.vertices, self.uvs, self.indices, self.meshes, self.models = [], [], [], [], []
        self.texture_paths = []
        self.mesh_count_total = 0
        self.textures_rgba8 = None  # (layers, h, w, 4) uint8

    # src/scene.cpp:110-120
    def add_material(self, mtype, color, extra):
        self.materials.append((int(mtype), [f32(c) for c in color], f32(extra)))

    def add_sphere(self, pos, r, mat):
        self.spheres.append(([f32(c) for c in pos], f32(r), int(mat)))

    def add_plane(self, pos, normal, mat):
        self.planes.append(([f32(c) for c in pos], [f32(c) for c in normal], int(mat)))

    # src/scene.cpp:122-143
    def add_lens(self, pos, normal, r1, r2, h, mat):
        pos = np.asarray(pos, dtype=f32)
        normal = np.asarray(normal, dtype=f32)
        r1, r2, h = f32(r1), f32(r2), f32(h)
        assert r1 >= h and r2 >= h
        t1 = _sqrtf(r1 * r1 - h * h)
        p1 = pos + normal * t1
        t2 = _sqrtf(r2 * r2 - h * h)
        p2 = pos - normal * t2
        self.lenses.append((pos, p1.astype(f32), p2.astype(f32), r1, r2, int(mat)))

    # src/scene.cpp:192-295
    def load_model(self, path, mat, transform=None, texture_resolver=None):
        transform = mat4_identity() if transform is None else transform
        if len(self.materials) <= mat:
            raise SceneError("ERROR: MATERIAL OF ID: %d DOES NOT EXIST" % mat)
        if not os.path.exists(path):
            raise SceneError("ERROR: Assimp: Unable to open file \"%s\"." % path)
        meshes = load_obj(path)
        for m in meshes:
            vertex_anchor = len(self.vertices)
            index_anchor = len(self.indices)
            for i in range(len(m["pos"])):
                self.vertices.append(transform_vertex(m["pos"][i], transform))
                if m["uv"] is not None:
                    self.uvs.append(m["uv"][i].astype(f32))
            self.indices += [int(i) for i in m["idx"]]
            tex_id = MISS
            if self.materials[mat][0] == T_TEXTURED:
                if m["texture"] is None:
                    raise SceneError("ERROR: MESH HAS NO TEXTURE APPLIED, USE A DIFFERENT MATERIAL")
                if m["texture"] in self.texture_paths:
                    tex_id = self.texture_paths.index(m["texture"])
                else:
                    tex_id = len(self.texture_paths)
                    self.texture_paths.append(m["texture"])
            self.meshes.append((vertex_anchor, index_anchor, len(m["idx"]) // 3, tex_id))
        self.models.append((self.mesh_count_total, len(meshes), int(mat)))
        self.mesh_count_total += len(meshes)

    # src/scene.cpp:145-190; `resolve` maps the .mtl's path string to a readable file
    def load_textures(self, resolve=None):
        if not self.models:
            self.textures_rgba8 = None
            return
        if not self.texture_paths:
            raise SceneError("ERROR: TEXTURE COUNT = 0")
        layers = []
        for i, p in enumerate(self.texture_paths):
            q = resolve(p) if resolve else p
            if not os.path.exists(q):
                raise SceneError("ERROR: STBimage: COULD NOT FIND THE TEXTURE")
            t = load_texture(q)
            if layers and t.shape != layers[0].shape:
                raise SceneError("ERROR: TEXTURES HAVE DIFFERENT SIZES: TEMPLATE: %d x %d, TEXTURE ID(%d): %d x %d"
                                 % (layers[0].shape[1], layers[0].shape[0], i, t.shape[1], t.shape[0]))
            layers.append(t)
        self.textures_rgba8 = np.stack(layers)

    # ---- packed arrays in the reference's host AoS layouts
    def arrays(self):
        def v4(v):
            return [float(v[0]), float(v[1]), float(v[2]), 0.0]
        mats = np.zeros(len(self.materials), dtype=MATERIAL_DT)
        for i, (t, c, e) in enumerate(self.materials):
            mats[i] = (t, v4(c), e)
        sph = np.zeros(len(self.spheres), dtype=SPHERE_DT)
        for i, (p, r, m) in enumerate(self.spheres):
            sph[i] = (v4(p), r, m)
        pla = np.zeros(len(self.planes), dtype=PLANE_DT)
        for i, (p, n, m) in enumerate(self.planes):
            pla[i] = (v4(p), v4(n), m)
        len_ = np.zeros(len(self.lenses), dtype=LENS_DT)
        for i, (p, p1, p2, r1, r2, m) in enumerate(self.lenses):
            len_[i] = (v4(p), v4(p1), v4(p2), r1, r2, m)
        verts = np.zeros((len(self.vertices), 4), dtype=f32)
        if self.vertices:
            verts[:, :3] = np.asarray(self.vertices, dtype=f32)
        uvs = np.asarray(self.uvs, dtype=f32).reshape(-1, 2)
        idx = np.asarray(self.indices, dtype=np.uint32)
        meshes = np.array(self.meshes, dtype=MESH_DT) if self.meshes else np.zeros(0, dtype=MESH_DT)
        models = np.array(self.models, dtype=MODEL_DT) if self.models else np.zeros(0, dtype=MODEL_DT)
        out = {"materials": mats, "spheres": sph, "planes": pla, "lenses": len_, "vertices": verts, "uvs": uvs,
               "indices": idx, "meshes": meshes, "models": models}
        if self.textures_rgba8 is not None:
            out["textures_rgba8"] = self.textures_rgba8
        return out


# ----------------------------------------------------------------------------- .scene parser
_RE_DELIM = re.compile(r",(?![^(]*\))")
_NUM = r"(?=(\d|\.|[-+]))([-+]?(\d*)+(\.\d+)?)"
_RE_VEC = re.compile(r"\s*(\(" + _NUM + r"\,\s*" + _NUM + r"\,\s*" + _NUM + r"\))\s*")
_RE_FLOAT = re.compile(r"\s*" + _NUM + r"\s*")
_RE_UINT = re.compile(r"\s*\d\s*")
_RE_PATH = re.compile(r"\s*\".*?\"\s*")


def _tokens(line):
    toks = _RE_DELIM.split(line)
    if len(toks) > 1 and toks[-1] == "":
        toks.pop()  # std::sregex_token_iterator(-1) yields no empty trailing suffix
    return toks


class _Iter:
    def __init__(self, toks):
        self.toks, self.i = toks, 0

    def next(self):
        if self.i >= len(self.toks):
            raise SceneError("ERROR: SCENE: NOT ENOUGH PARAMETERS")
        w = self.toks[self.i]
        return w

    def advance(self):
        self.i += 1


def _stof(s):
    m = re.match(r"\s*[-+]?(\d+\.?\d*|\.\d+)", s)
    if not m:
        raise ValueError("stof")
    return f32(float(m.group(0)))


def _get_vec(it):
    w = it.next()
    if not _RE_VEC.fullmatch(w):
        raise SceneError("ERROR: SCENE: IMPROPER VECTOR: " + w)
    w = w[w.find("(") + 1:]
    p = w.find(",")
    x = _stof(w[:p]); w = w[p + 1:]
    p = w.find(",")
    y = _stof(w[:p]); w = w[p + 1:]
    p = w.find(")")
    z = _stof(w[:p])
    it.advance()
    return [x, y, z]


def _get_float(it):
    w = it.next()
    if not _RE_FLOAT.fullmatch(w):
        raise SceneError("ERROR: SCENE: IMPROPER FLOAT: " + w)
    it.advance()
    return _stof(w)


def _get_uint(it):
    w = it.next()
    if not _RE_UINT.fullmatch(w):
        raise SceneError("ERROR: SCENE: IMPROPER UNSIGNED INT: " + w)
    it.advance()
    return int(w)


def _get_path(it):
    w = it.next()
    if not _RE_PATH.fullmatch(w):
        raise SceneError("ERROR: SCENE: IMPROPER PATH: " + w)
    b, e = w.find('"'), w.rfind('"')
    it.advance()
    return w[b + 1:b + 1 + max(e - 2, 0)]  # substr(pos_begin + 1, pos_end - 2)


def parse_scene_file(path, root=".", texture_resolver=None):
    """Returns a Scene.  Model paths are resolved relative to `root` (the reference runs with the
    repository root as working directory)."""
    try:
        with open(path, "r") as fh:
            data = fh.read()
    except OSError as e:
        raise SceneError("ERROR: SCENE: NOT SUCCESFULLY READ: " + str(e))
    sc = Scene()
    mode = None
    model = mat4_identity()
    lines = data.split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    for line in lines:
        p = line.find("#")
        if p >= 0:
            line = line[:p]
        if line == "":
            continue
        p = line.find(":")
        if p >= 0:
            word = line[:p]
            if word in ("MATERIALS", "SPHERES", "PLANES", "LENSES", "MODELS"):
                mode = word
                continue
            if mode == "MODELS":
                it = _Iter(_tokens(line[p + 1:]))
                if word == "translate":
                    model = mat4_translate(model, _get_vec(it))
                elif word == "rotate":
                    ang = glm_radians(_get_float(it))
                    model = mat4_rotate(model, ang, _get_vec(it))
                elif word == "scale":
                    model = mat4_scale(model, _get_vec(it))
                elif word == "load":
                    mp = _get_path(it)
                    mid = _get_uint(it)
                    sc.load_model(os.path.join(root, mp), mid, model)
                    model = mat4_identity()
            else:
                raise SceneError("ERROR: SCENE: OPERATION " + word + " DOES NOT EXIST")
        else:
            it = _Iter(_tokens(line))
            if mode == "MATERIALS":
                word = it.next()
                if word not in MAT_NAMES:
                    raise SceneError("ERROR: SCENE: MATERIAL: " + word + " DOES NOT EXIST")
                it.advance()
                c = _get_vec(it); e = _get_float(it)
                sc.add_material(MAT_NAMES[word], c, e)
            elif mode == "SPHERES":
                p3 = _get_vec(it); r = _get_float(it); m = _get_uint(it)
                sc.add_sphere(p3, r, m)
            elif mode == "PLANES":
                p3 = _get_vec(it); n3 = _get_vec(it); m = _get_uint(it)
                sc.add_plane(p3, n3, m)
            elif mode == "LENSES":
                p3 = _get_vec(it); n3 = _get_vec(it)
                r1 = _get_float(it); r2 = _get_float(it); h = _get_float(it); m = _get_uint(it)
                sc.add_lens(p3, n3, r1, r2, h, m)
            else:
                raise SceneError("ERROR: SCENE: OPERATION NOT SPECIFIED")
    return sc


# ----------------------------------------------------------------------------- camera
def camera_data(fov_deg, aspect, pos=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0):
    """Camera::Camera + updateVectors + transferData (src/camera.cpp:26-51,94-118): 12 floats
    [pos, lower_left_corner, horizontal, vertical]."""
    fov = f32(int(fov_deg))  # constructor takes int camera_fov
    aspect = f32(aspect)
    # `fov*M_PI/180.0f` is evaluated in double (M_PI is a double) and stored to a float
    angle = f32(np.float64(fov) * np.float64(np.pi) / np.float64(f32(180.0)))
    # `tan(angle*0.5f)`: float overload (libc++ on the author's platform exposes it globally)
    half_height = _tanf(angle * f32(0.5))
    half_width = aspect * half_height
    rp, ry = glm_radians(f32(pitch)), glm_radians(f32(yaw))
    w = np.array([_cosf(rp) * _sinf(ry), _sinf(rp), _cosf(rp) * _cosf(ry)], dtype=f32)
    w = _glm_normalize(w)
    up = np.array([0.0, -1.0, 0.0], dtype=f32)
    u = _glm_normalize(_cross(w, up))
    v = _cross(u, w)
    llc = w - (half_width * u + half_height * v)
    horizontal = f32(2.0) * half_width * u
    vertical = f32(2.0) * half_height * v
    return np.concatenate([np.asarray(pos, dtype=f32), llc, horizontal, vertical]).astype(f32)


class CameraPy:
    """The reference's Camera with its controls (src/camera.cpp:16-118), float for float: move / rotate / zoom, the two
    speed switches, setSize, and the two spellings of degrees -> radians (constructor: fov*M_PI/180.0f, setFov:
    fov*M_PI*0.0055556f, both evaluated in double and stored to a float)."""
    SLOW, NORMAL, FAST = f32(0.3), f32(1.0), f32(5.0)          # :16-18
    SENS, ZOOM_MIN, ZOOM_MAX, ZOOM_SPEED = f32(0.2), f32(90.0), f32(10.0), f32(0.5)   # :19-22

    def __init__(self, fov_deg, aspect, pos=(0.0, 0.0, 0.0), yaw=0.0, pitch=0.0):     # :46-51
        self.fov, self.aspect = f32(int(fov_deg)), f32(aspect)
        self.pos = np.asarray(pos, dtype=f32).copy()
        self.yaw, self.pitch, self.speed = f32(yaw), f32(pitch), self.SLOW
        angle = f32(np.float64(self.fov) * np.float64(np.pi) / np.float64(f32(180.0)))
        self.half_height = _tanf(angle * f32(0.5))
        self.half_width = self.aspect * self.half_height
        self._update()

    def _update(self):                                          # updateVectors, :26-37
        rp, ry = glm_radians(self.pitch), glm_radians(self.yaw)
        w = np.array([_cosf(rp) * _sinf(ry), _sinf(rp), _cosf(rp) * _cosf(ry)], dtype=f32)
        self.w = _glm_normalize(w)
        self.u = _glm_normalize(_cross(self.w, np.array([0.0, -1.0, 0.0], dtype=f32)))
        self.v = _cross(self.u, self.w)

    def _set_fov(self):                                         # :39-44
        angle = f32(np.float64(self.fov) * np.float64(np.pi) * np.float64(f32(0.0055556)))
        self.half_height = _tanf(angle * f32(0.5))
        self.half_width = self.aspect * self.half_height
        self._update()

    def move(self, direction, dt):                              # :53-59; 0 forward, 1 back, 2 left, 3 right
        ds = self.speed * f32(dt)
        step = ((self.w if direction < 2 else self.u) * ds).astype(f32)
        self.pos = (self.pos + step if direction in (0, 3) else self.pos - step).astype(f32)

    def rotate(self, x, y):                                     # :61-70
        self.yaw = f32(self.yaw + f32(x) * self.SENS * self.fov / self.ZOOM_MAX)
        self.pitch = f32(self.pitch + f32(y) * self.SENS * self.fov / self.ZOOM_MAX)
        self.pitch = min(max(self.pitch, f32(-89.0)), f32(89.0))
        self.yaw = f32(math.fmod(float(self.yaw), 360.0))       # fmod of floats is exact
        self._update()

    def zoom(self, scroll):                                     # :72-77
        self.fov = f32(self.fov + f32(scroll) * self.ZOOM_SPEED)
        self.fov = min(max(self.fov, self.ZOOM_MAX), self.ZOOM_MIN)
        self._set_fov()

    def set_faster_speed(self, on): self.speed = self.FAST if on else self.NORMAL     # :79-82
    def set_slower_speed(self, on): self.speed = self.SLOW if on else self.NORMAL     # :84-87

    def set_size(self, aspect):                                 # :89-92
        self.aspect = f32(aspect)
        self._set_fov()

    def data(self):                                             # transferData, :94-118 (+ the tail of updateVectors)
        llc = self.w - (self.half_width * self.u + self.half_height * self.v)
        horizontal = f32(2.0) * self.half_width * self.u
        vertical = f32(2.0) * self.half_height * self.v
        return np.concatenate([self.pos, llc, horizontal, vertical]).astype(f32)


def _cross(a, b):
    return np.array([a[1] * b[2] - b[1] * a[2], a[2] * b[0] - b[2] * a[0], a[0] * b[1] - b[0] * a[1]], dtype=f32)


def _glm_normalize(v):
    inv = f32(1.0) / _sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
    return (v * inv).astype(f32)
