import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "opencl-raytracing_b200"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    """The plain-C oracle (oracle/rt_oracle.c); built on demand."""
    import oracle_py
    lib = os.path.join(ROOT, "oracle", "lib", "liboracle_port.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"])
    return oracle_py.PortOracle()


@pytest.fixture(scope="session")
def ref():
    """The reference's own kernel compiled as C++ (oracle/_ref); prebuilt where /root/reference exists."""
    import oracle_py
    try:
        return oracle_py.RefOracle("strict")
    except (FileNotFoundError, OSError) as e:
        pytest.skip("oracle/_ref not built: %s" % e)


@pytest.fixture(scope="session")
def scenes():
    """name -> (arrays, float textures) from the committed fixtures."""
    from rtb200 import fixtures as fx
    out = {}
    for name, loader in (("default", fx.default_scene), ("c3", fx.c3_scene)):
        arr, tex8 = loader()
        out[name] = (arr, fx.textures_to_float(tex8))
    return out


@pytest.fixture(scope="session")
def golden():
    return {n: np.load(os.path.join(GOLDEN, "golden_%s.npz" % n)) for n in ("default", "c3")}


@pytest.fixture(scope="session")
def table(port, golden):
    import hashlib
    t = port.make_table(int(golden["default"]["table_seed"]))
    assert hashlib.sha256(t.tobytes()).hexdigest() == str(golden["default"]["table_sha256"])
    return t
