"""GPU test of the drop-in C++ surface: the headless driver (host/rt_headless.cpp) builds Camera + RayTracer exactly as
the reference's main() does and runs render() once then renderAgain() (main.cpp:85-96).  Its image must be bit-exact
against the oracle fed with the same scene file, table seed and camera."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_py as op

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ASSETS = os.path.join(ROOT, "tests", "golden", "assets")
EXE = os.path.join(ROOT, "opencl-raytracing_b200", "bin", "rt_headless")


def run(args, tmp):
    raw = str(tmp / "img.f32")
    out = subprocess.run([EXE, "--scene", "shapes.scene", "--root", ASSETS, "--textures", ASSETS, "--raw", raw,
                          "--out", str(tmp / "img.tga")] + args, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    return json.loads(out.stdout.strip().splitlines()[-1]), raw


def test_headless_driver_matches_oracle(port, tmp_path):
    import rtb200
    W, H, spp, depth = 200, 150, 5, 12
    info, raw = run(["--width", str(W), "--height", str(H), "--spp", str(spp), "--depth", str(depth), "--arith", "strict",
                     "--rng", "table", "--table-seed", "777"], tmp_path)
    img = np.fromfile(raw, np.float32).reshape(H, W, 4)
    arr, tex = rtb200.load_scene_file("shapes.scene", ASSETS, ASSETS)
    cam = rtb200.camera_data(60, W / H)
    table = port.make_table(777)
    want, rays, _ = port.render(op.make_desc(arr), tex, cam, table, W, H, 0, spp, port.params(max_depth=depth))
    assert info["frames"] == spp and info["samples_in_image"] == spp
    assert int((img.view(np.uint32) != want.view(np.uint32)).any(-1).sum()) <= 2
    assert abs(info["rays"] - rays) <= 40
    assert os.path.getsize(str(tmp_path / "img.tga")) == 18 + W * H * 3


def test_headless_script_resets_on_camera_motion(port, tmp_path):
    """A camera move restarts the accumulation with render() (sample 0), as main.cpp:91-94 does."""
    import rtb200
    W, H = 96, 64
    script = tmp_path / "s.txt"
    script.write_text("frames 3\nmove forward 2.0\nrotate 10 -5\nframes 2\n")
    info, raw = run(["--width", str(W), "--height", str(H), "--arith", "strict", "--table-seed", "5", "--script", str(script)], tmp_path)
    assert info["frames"] == 5 and info["samples_in_image"] == 2
    img = np.fromfile(raw, np.float32).reshape(H, W, 4)
    # Camera starts slow (0.3): forward 2.0 s moves 0.6 along +z; rotate(10,-5) at fov 60: yaw += 10*0.2*60/10, pitch -= 5*0.2*6
    cam = rtb200.camera_data(60, W / H, (0.0, 0.0, np.float32(0.3) * np.float32(2.0)), 12.0, -6.0)
    arr, tex = rtb200.load_scene_file("shapes.scene", ASSETS, ASSETS)
    want, _, _ = port.render(op.make_desc(arr), tex, cam, port.make_table(5), W, H, 0, 2, port.params(max_depth=30))
    assert int((img.view(np.uint32) != want.view(np.uint32)).any(-1).sum()) <= 2


def test_headless_reports_scene_errors_like_the_reference(tmp_path):
    out = subprocess.run([EXE, "--scene", "bad_material.scene", "--root", ASSETS], capture_output=True, text=True, timeout=120)
    assert out.returncode != 0 and "ERROR: SCENE: MATERIAL: shiny DOES NOT EXIST" in out.stderr


def test_headless_checkpoint_resume_is_bit_identical(tmp_path):
    """3 samples + checkpoint, then resume + 4 more samples == 7 samples uninterrupted (the sqrt(mean) chain of
    kernels/raytracer.cl:524-531 continues from the restored image and sample counter)."""
    W, H = 120, 80
    common = ["--width", str(W), "--height", str(H), "--arith", "strict", "--rng", "table", "--table-seed", "9", "--depth", "10"]
    ck = str(tmp_path / "state.rtck")
    a_dir, b_dir, c_dir = tmp_path / "a", tmp_path / "b", tmp_path / "c"
    for d in (a_dir, b_dir, c_dir):
        d.mkdir()
    info_a, _ = run(common + ["--spp", "3", "--checkpoint", ck], a_dir)
    assert info_a["samples_in_image"] == 3 and os.path.getsize(ck) == 24 + W * H * 16      # "RTC2", w, h, samples, u64 fingerprint
    info_b, raw_b = run(common + ["--spp", "4", "--resume", ck], b_dir)
    info_c, raw_c = run(common + ["--spp", "7"], c_dir)
    assert info_b["samples_in_image"] == 7 and info_c["samples_in_image"] == 7
    assert np.array_equal(np.fromfile(raw_b, np.uint32), np.fromfile(raw_c, np.uint32))
    bad = subprocess.run([EXE, "--scene", "shapes.scene", "--root", ASSETS, "--textures", ASSETS, "--width", "64", "--height", "64",
                          "--resume", ck], capture_output=True, text=True, timeout=120)
    assert bad.returncode == 2 and "cannot resume" in bad.stderr
    # a checkpoint continued with another camera, depth, arithmetic / RNG mode or table seed is refused, not blended
    for change in (["--depth", "11"], ["--table-seed", "10"], ["--yaw", "3"], ["--arith", "fast"]):
        args = list(common)
        if change[0] in args:
            args[args.index(change[0]) + 1] = change[1]
        else:
            args += change
        r = subprocess.run([EXE, "--scene", "shapes.scene", "--root", ASSETS, "--textures", ASSETS] + args + ["--spp", "1", "--resume", ck],
                           capture_output=True, text=True, timeout=120)
        assert r.returncode == 2 and "different camera, depth" in r.stderr, (change, r.stderr)
