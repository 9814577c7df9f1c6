"""CPU tests of the C++ host classes (SceneCreator, Camera — include/scene.h, include/camera.h) through their C view
(rt_scene_load, rt_camera_data): compared bit for bit with the oracle's numpy restatement of the reference's host
code (oracle/scene_py.py) on repository-owned assets, plus the reference's error messages (src/scene.cpp)."""
import os

import numpy as np
import pytest

import scene_py as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ASSETS = os.path.join(ROOT, "tests", "golden", "assets")
FIELDS = {"materials": (sp.MATERIAL_DT, ["type", "color", "extra"]), "spheres": (sp.SPHERE_DT, ["pos", "r", "mat"]),
          "planes": (sp.PLANE_DT, ["pos", "normal", "mat"]), "lenses": (sp.LENS_DT, ["pos", "p1", "p2", "r1", "r2", "mat"])}


def assert_same_scene(got, want):
    for name, (dt, fields) in FIELDS.items():
        g = got[name].reshape(-1).view(dt)
        w = want[name]
        assert g.shape == w.shape, name
        if len(g) == 0:
            continue
        for f in fields:
            a, b = np.ascontiguousarray(g[f]), np.ascontiguousarray(w[f])
            if a.dtype.kind == "f":
                a, b = a.reshape(len(g), -1)[:, :3 if a.ndim > 1 else None], b.reshape(len(w), -1)[:, :3 if b.ndim > 1 else None]
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (name, f)
            else:
                assert np.array_equal(a, b), (name, f)
    assert np.array_equal(got["vertices"].view(np.float32)[:, :3].view(np.uint32), want["vertices"][:, :3].view(np.uint32))
    assert np.array_equal(got["uvs"].view(np.float32).view(np.uint32), np.ascontiguousarray(want["uvs"]).view(np.uint32))
    assert np.array_equal(got["indices"].view(np.uint32).ravel(), want["indices"])
    assert np.array_equal(got["meshes"].view(np.uint32), np.ascontiguousarray(want["meshes"]).view(np.uint32).reshape(-1, 4))
    assert np.array_equal(got["models"].view(np.uint32), np.ascontiguousarray(want["models"]).view(np.uint32).reshape(-1, 3))


def test_own_scene_matches_python_restatement():
    import rtb200
    arr, tex = rtb200.load_scene_file("shapes.scene", ASSETS, ASSETS)
    sc = sp.parse_scene_file(os.path.join(ASSETS, "shapes.scene"), root=ASSETS)
    sc.load_textures(lambda p: os.path.join(ASSETS, os.path.basename(p)))
    want = sc.arrays()
    assert_same_scene(arr, want)
    assert [len(arr[k]) for k in ("materials", "spheres", "planes", "lenses", "models", "meshes")] == [6, 4, 1, 1, 1, 2]
    assert len(arr["vertices"]) == 4 * 3 + 3 * 2 + 4 and len(arr["indices"]) == 3 * (2 * 3 + 2 + 2)   # corners unrolled, quads fanned
    assert tex.shape == (2, 32, 32, 4)
    assert np.array_equal(tex.view(np.uint32), sp.texture_from_rgba8(want["textures_rgba8"]).view(np.uint32))


@pytest.mark.skipif(not os.path.exists("/root/reference/assets/scenes/scene.scene"), reason="reference tree not present")
def test_reference_default_scene_matches_python_restatement_and_fixture(scenes):
    import rtb200
    arr, tex = rtb200.load_scene_file("assets/scenes/scene.scene", "/root/reference")
    sc = sp.parse_scene_file("/root/reference/assets/scenes/scene.scene", root="/root/reference")
    sc.load_textures(lambda p: "/root/reference/assets/textures/" + os.path.basename(p))
    assert_same_scene(arr, sc.arrays())
    assert [len(arr[k]) for k in ("materials", "spheres", "planes", "lenses", "models", "meshes")] == [9, 8, 1, 1, 2, 2]
    assert len(arr["vertices"]) == 48 and len(arr["indices"]) == 72
    assert np.array_equal(tex.view(np.uint32), scenes["default"][1].view(np.uint32))   # the committed fixture


@pytest.mark.parametrize("name,message", [
    ("bad_material.scene", "ERROR: SCENE: MATERIAL: shiny DOES NOT EXIST"),
    ("bad_vector.scene", "ERROR: SCENE: IMPROPER VECTOR: (0 , 0, 3)"),
    ("bad_uint.scene", "ERROR: SCENE: IMPROPER UNSIGNED INT:  10"),
    ("bad_float.scene", "ERROR: SCENE: IMPROPER FLOAT:  1e3"),
    ("missing_param.scene", "ERROR: SCENE: NOT ENOUGH PARAMETERS"),
    ("no_section.scene", "ERROR: SCENE: OPERATION NOT SPECIFIED"),
    ("bad_operation.scene", "ERROR: SCENE: OPERATION rotate DOES NOT EXIST"),
    ("blank_line.scene", "ERROR: SCENE: MATERIAL:     DOES NOT EXIST"),
    ("bad_path.scene", "ERROR: SCENE: IMPROPER PATH:  shapes.obj"),
    ("no_material.scene", "ERROR: MATERIAL OF ID: 3 DOES NOT EXIST"),
    ("untextured_model.scene", "ERROR: TEXTURE COUNT = 0"),
    ("does_not_exist.scene", "ERROR: SCENE: NOT SUCCESFULLY READ"),
])
def test_error_messages_of_the_reference(name, message):
    import rtb200
    with pytest.raises(rtb200.RtError) as e:
        rtb200.load_scene_file(name, ASSETS, ASSETS)
    assert message in str(e.value)
    if name != "does_not_exist.scene":
        with pytest.raises(sp.SceneError) as e2:
            s = sp.parse_scene_file(os.path.join(ASSETS, name), root=ASSETS)
            s.load_textures(lambda p: os.path.join(ASSETS, os.path.basename(p)))
        assert message in str(e2.value)


def test_camera_matches_python_restatement():
    import rtb200
    for args in ((60, 800 / 600), (60, 1920 / 1080, (0, 0, 0), 0, 0), (45, 1.5, (1, -2, 3), 30.0, -10.0), (90, 1.0, (0, 0, 0), 359.0, 89.0)):
        assert np.array_equal(rtb200.camera_data(*args).view(np.uint32), sp.camera_data(*args).view(np.uint32))
    c = rtb200.camera_data(60, 4 / 3)
    assert np.allclose(c[3:6] + 0.5 * c[6:9] + 0.5 * c[9:12], [0, 0, 1], atol=1e-6)      # centre of the view looks down +z
    assert c[10] < 0                                                                        # screen-up is world -y


def test_tga_layout(tmp_path):
    import rtb200
    img = np.zeros((2, 3, 4), np.float32)
    img[0, 0, :3] = (1.0, 0.5, 0.0)
    img[1, 2, :3] = (0.0, 0.0, 2.0)    # clamped
    p = str(tmp_path / "t.tga")
    assert rtb200.lib().rt_write_tga(p.encode(), img.ctypes.data, 3, 2) == 0
    b = open(p, "rb").read()
    assert len(b) == 18 + 3 * 2 * 3
    assert np.frombuffer(b[:18], np.uint16).tolist() == [0, 2, 0, 0, 0, 0, 3, 2, 24]
    assert list(b[18:21]) == [0, 128, 255] and list(b[18 + 15:18 + 18]) == [255, 0, 0]    # BGR, bottom row first


def _random_scene_text(rng):
    """A random scene in the reference's grammar (src/scene.cpp:297-461): every section, comments, odd but legal number
    spellings, several models with a running transform."""
    def num(lo, hi):
        v = rng.uniform(lo, hi)
        s = rng.choice(["%.3f", "%.1f", "%.6f", "%d"])
        t = s % (int(v) if s == "%d" else v)
        if t.startswith("0.") and rng.random() < 0.3:
            t = t[1:]                                   # ".4"
        if not t.startswith("-") and rng.random() < 0.2:
            t = "+" + t                                 # "+1.1"
        return t
    def vec(lo, hi):
        return "(%s, %s, %s)" % (num(lo, hi), num(lo, hi), num(lo, hi))
    types = ["refractive", "reflective", "dielectric", "diffuse", "light"]
    lines = ["# random scene", "MATERIALS:", "textured, (1, 1, 1), 1   #0"]
    n_mat = int(rng.integers(2, 10))
    for i in range(1, n_mat):
        lines.append("%s, %s, %s%s" % (rng.choice(types), vec(0, 1), num(0, 2), "  # m%d" % i if rng.random() < 0.5 else ""))
    lines += ["", "SPHERES:"]
    for _ in range(int(rng.integers(0, 6))):
        lines.append("%s, %s, %d" % (vec(-5, 5), num(0.1, 3), rng.integers(1, n_mat)))
    lines += ["PLANES:"]
    for _ in range(int(rng.integers(0, 3))):
        lines.append("%s, %s, %d" % (vec(-5, 5), vec(-1, 1), rng.integers(1, n_mat)))
    lines += ["# lenses", "LENSES:"]
    for _ in range(int(rng.integers(0, 3))):
        lines.append("%s, %s, %s, %s, %s, %d" % (vec(-5, 5), vec(-1, 1), num(2, 9), num(2, 9), num(0.2, 1.5), rng.integers(1, n_mat)))
    lines += ["MODELS:"]
    for m in range(int(rng.integers(1, 4))):
        for _ in range(int(rng.integers(0, 4))):
            op = rng.choice(["translate", "rotate", "scale"])
            if op == "rotate":
                lines.append("rotate: %s, %s" % (num(-180, 180), vec(0.1, 1)))
            else:
                lines.append("%s: %s" % (op, vec(0.2, 3)))
        lines.append('load: "shapes.obj", %d' % (0 if m == 0 else rng.integers(0, n_mat)))   # a scene without a textured model is an error (src/scene.cpp:181-190)
        if rng.random() < 0.5:
            lines.append("")
    return "\n".join(lines) + "\n"


@pytest.mark.parametrize("seed", range(12))
def test_random_scene_files_match_python_restatement(tmp_path, seed):
    """The C++ SceneCreator (parser, running model matrix, OBJ import, lens precompute) against the numpy restatement of
    src/scene.cpp on seeded random scene files: every array bit for bit."""
    import shutil
    import rtb200
    rng = np.random.default_rng(1000 + seed)
    for f in ("shapes.obj", "shapes.mtl", "tex_a.png", "tex_b.png"):
        shutil.copy(os.path.join(ASSETS, f), tmp_path / f)
    (tmp_path / "r.scene").write_text(_random_scene_text(rng))
    arr, tex = rtb200.load_scene_file("r.scene", str(tmp_path), str(tmp_path))
    sc = sp.parse_scene_file(str(tmp_path / "r.scene"), root=str(tmp_path))
    sc.load_textures(lambda p: str(tmp_path / os.path.basename(p)))
    want = sc.arrays()
    assert_same_scene(arr, want)
    assert np.array_equal(tex.view(np.uint32), sp.texture_from_rgba8(want["textures_rgba8"]).view(np.uint32))


@pytest.mark.parametrize("seed", range(8))
def test_camera_controls_match_python_restatement(seed):
    """Camera::move / rotate / zoom / setFasterSpeed / setSlowerSpeed / setSize (src/camera.cpp:53-92) through
    rt_camera_run against the numpy restatement, bit for bit, on seeded random control sequences — including the pitch
    clamp at +-89, yaw wrapping at 360, the fov limits 10..90 and setFov's 0.0055556 spelling of pi/180."""
    import rtb200
    rng = np.random.default_rng(77 + seed)
    fov, aspect = int(rng.integers(20, 90)), float(np.float32(rng.uniform(0.8, 2.4)))
    pos = rng.uniform(-3, 3, 3).astype(np.float32)
    yaw, pitch = float(np.float32(rng.uniform(-200, 400))), float(np.float32(rng.uniform(-80, 80)))
    ops = []
    for _ in range(int(rng.integers(5, 60))):
        code = int(rng.integers(0, 9))
        a = float(np.float32(rng.uniform(-40, 40) if code == 5 else rng.uniform(-300, 300) if code == 4 else rng.uniform(0.0, 2.5)))
        b = float(np.float32(rng.uniform(-300, 300)))
        if code in (6, 7):
            a = float(rng.integers(0, 2))
        ops.append((code, a, b))
    cam = sp.CameraPy(fov, aspect, pos, yaw, pitch)
    for code, a, b in ops:
        if code < 4: cam.move(code, a)
        elif code == 4: cam.rotate(a, b)
        elif code == 5: cam.zoom(a)
        elif code == 6: cam.set_faster_speed(a != 0)
        elif code == 7: cam.set_slower_speed(a != 0)
        else: cam.set_size(a)
    got = rtb200.camera_run(fov, aspect, pos, yaw, pitch, ops)
    assert np.array_equal(got.view(np.uint32), cam.data().view(np.uint32)), (got, cam.data())
    assert np.array_equal(rtb200.camera_run(fov, aspect, pos, yaw, pitch, []).view(np.uint32),
                          rtb200.camera_data(fov, aspect, pos, yaw, pitch).view(np.uint32))
    with pytest.raises(rtb200.RtError):
        rtb200.camera_run(fov, aspect, pos, yaw, pitch, [(9, 0, 0)])


def _write_png_rgba(path, img, filter_type):
    """Minimal PNG writer (8-bit RGBA, non-interlaced) that uses ONE filter type for every row — the files image tools
    write mix them adaptively, so this is how each reconstruction rule of the reader (PNG spec 9.2) gets its own test."""
    import struct
    import zlib
    h, w, _ = img.shape
    raw = bytearray()
    prev = np.zeros((w * 4,), np.int32)
    for y in range(h):
        cur = img[y].reshape(-1).astype(np.int32)
        left = np.concatenate([np.zeros(4, np.int32), cur[:-4]])
        upleft = np.concatenate([np.zeros(4, np.int32), prev[:-4]])
        if filter_type == 0: pred = np.zeros_like(cur)
        elif filter_type == 1: pred = left
        elif filter_type == 2: pred = prev
        elif filter_type == 3: pred = (left + prev) >> 1
        else:
            p = left + prev - upleft
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - upleft)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, upleft))
        raw.append(filter_type)
        raw += bytes(((cur - pred) & 255).astype(np.uint8))
        prev = cur
    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    comp = zlib.compress(bytes(raw), 6)
    half = len(comp) // 2                       # two IDAT chunks: the reader must concatenate them
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"tEXt", b"Comment\x00filter %d" % filter_type) + chunk(b"IDAT", comp[:half]) + chunk(b"IDAT", comp[half:]) +
                chunk(b"IEND", b""))


@pytest.mark.parametrize("filter_type", range(5))
def test_png_reader_every_filter_type(tmp_path, filter_type):
    """The built-in PNG reader behind SceneCreator::loadTextures (the reference calls stbi_loadf, src/scene.cpp:158): every
    row filter of the PNG spec, split IDAT, an ancillary chunk; texels = stb's float conversion of the decoded bytes."""
    import shutil
    import rtb200
    rng = np.random.default_rng(5 + filter_type)
    for f in ("shapes.obj", "shapes.mtl", "shapes.scene"):
        shutil.copy(os.path.join(ASSETS, f), tmp_path / f)
    imgs = []
    for name in ("tex_a.png", "tex_b.png"):
        img = rng.integers(0, 256, (32, 32, 4), dtype=np.uint8)
        img[::3] = img[::3] // 64 * 64                      # some smooth rows, so that predictions matter
        _write_png_rgba(str(tmp_path / name), img, filter_type)
        imgs.append(img)
    arr, tex = rtb200.load_scene_file("shapes.scene", str(tmp_path), str(tmp_path))
    paths = [os.path.basename(p) for p in _texture_order(tmp_path)]
    want = sp.texture_from_rgba8(np.stack([imgs[("tex_a.png", "tex_b.png").index(p)] for p in paths]))
    assert tex.shape == want.shape and np.array_equal(tex.view(np.uint32), want.view(np.uint32))


def _texture_order(tmp_path):
    sc = sp.parse_scene_file(str(tmp_path / "shapes.scene"), root=str(tmp_path))
    return sc.texture_paths


@pytest.mark.parametrize("seed", range(10))
def test_random_obj_files_match_python_restatement(tmp_path, seed):
    """The OBJ / MTL import behind SceneCreator::loadModel (the reference goes through Assimp with Triangulate | FlipUVs,
    src/scene.cpp:192-295) on seeded random OBJ files: polygons of 3-6 corners (fan triangulation), positive and negative
    indices, v/vt and v/vt/vn corners, several objects and materials, under a random model transform — against the numpy
    restatement, every array bit for bit."""
    import shutil
    import rtb200
    rng = np.random.default_rng(300 + seed)
    for f in ("tex_a.png", "tex_b.png"):
        shutil.copy(os.path.join(ASSETS, f), tmp_path / f)
    (tmp_path / "m.mtl").write_text("newmtl first\nKd 1 1 1\nmap_Kd tex_a.png\n\nnewmtl second\nKd 1 1 1\nmap_Kd some/dir/tex_b.png\n")
    lines, nv, nt, nn = ["mtllib m.mtl"], 0, 0, 0
    for o in range(int(rng.integers(1, 4))):
        lines.append("o part%d" % o)
        v_new, t_new = int(rng.integers(4, 12)), int(rng.integers(3, 8))
        for _ in range(v_new):
            lines.append("v %.4f %.4f %.4f" % tuple(rng.uniform(-2, 2, 3)))
        for _ in range(t_new):
            lines.append("vt %.4f %.4f" % tuple(rng.uniform(0, 1, 2)))
        with_normals = rng.random() < 0.5
        if with_normals:
            for _ in range(3):
                lines.append("vn %.4f %.4f %.4f" % tuple(rng.uniform(-1, 1, 3)))
            nn += 3
        nv += v_new; nt += t_new
        lines.append("usemtl %s" % ("first" if o % 2 == 0 else "second"))
        for _ in range(int(rng.integers(1, 7))):
            corners = []
            for vi in rng.choice(np.arange(nv - v_new, nv), size=int(rng.integers(3, 7)) if v_new >= 6 else 3, replace=False):
                ti = int(rng.integers(nt - t_new, nt))
                if rng.random() < 0.3:
                    vs, ts = str(int(vi) - nv), str(ti - nt)                # relative (negative) indices
                else:
                    vs, ts = str(int(vi) + 1), str(ti + 1)
                corners.append("%s/%s/%d" % (vs, ts, int(rng.integers(nn - 3, nn)) + 1) if with_normals else "%s/%s" % (vs, ts))
            lines.append("f " + " ".join(corners))
    (tmp_path / "m.obj").write_text("\n".join(lines) + "\n")
    (tmp_path / "m.scene").write_text("MATERIALS:\ntextured, (1, 1, 1), 1\nMODELS:\nrotate: %.2f, (%.2f, %.2f, 1)\ntranslate: (%.2f, 0.5, 2)\n"
                                      "load: \"m.obj\", 0\n" % (rng.uniform(-90, 90), rng.uniform(0, 1), rng.uniform(0, 1), rng.uniform(-2, 2)))
    arr, tex = rtb200.load_scene_file("m.scene", str(tmp_path), str(tmp_path))
    sc = sp.parse_scene_file(str(tmp_path / "m.scene"), root=str(tmp_path))
    sc.load_textures(lambda p: str(tmp_path / os.path.basename(p)))
    want = sc.arrays()
    assert_same_scene(arr, want)
    assert len(arr["indices"]) % 3 == 0 and len(arr["indices"]) > 0
    assert np.array_equal(tex.view(np.uint32), sp.texture_from_rgba8(want["textures_rgba8"]).view(np.uint32))
