"""CPU test: the C-ABI library loads and exports every symbol include/rt_capi.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "rt_capi.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_reference_surface():
    names = declared_symbols()
    for must in ("rt_create", "rt_destroy", "rt_upload_scene", "rt_upload_textures", "rt_set_random_table",
                 "rt_render", "rt_render_again", "rt_render_samples", "rt_read_image", "rt_read_primary_ids"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import rtb200
    if not os.path.exists(rtb200.LIB_PATH):
        pytest.fail("librtb200.so is not built (run __graft_entry__.build())")
    lib = ctypes.CDLL(rtb200.LIB_PATH)
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing
    rtb200.lib()          # the binding resolves all its prototypes
    lib.rt_version.restype = ctypes.c_char_p
    assert b"rtb200" in lib.rt_version()


def test_no_cpu_fallback_without_a_device():
    """On a box without CUDA the product must fail loudly, not fall back."""
    import rtb200
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(rtb200.RtError) as e:
        rtb200.Context(64, 64)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_make_random_table_is_the_reference_loop(port):
    """rt_make_random_table (host helper in the product) equals the oracle's restatement of src/raytracer.cpp:67-93."""
    import numpy as np
    import rtb200
    t = rtb200.make_random_table(12345)
    assert np.array_equal(t.view(np.uint32), port.make_table(12345).view(np.uint32))
