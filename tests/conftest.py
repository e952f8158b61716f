import gzip
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN_DIR, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def bucky():
    with gzip.open(os.path.join(GOLDEN_DIR, "bucky32.u8.gz"), "rb") as f:
        return np.frombuffer(f.read(), dtype=np.uint8).reshape(32, 32, 32).copy()


@pytest.fixture(scope="session")
def port():
    import oracle
    return oracle.Port()


@pytest.fixture(scope="session")
def ref():
    import oracle
    if not oracle.ref_available():
        pytest.skip("oracle/_ref/libvolr_ref.so not built (needs /root/reference)")
    return oracle.Ref()


@pytest.fixture(scope="session")
def R():
    """volume-rendering_b200.renderer (ctypes over libvr5.so); builds the library if it is missing."""
    import vr5_loader
    pkg = vr5_loader.load()
    if not os.path.exists(pkg.LIB_PATH):
        pkg.build()
    return vr5_loader.renderer_module()


@pytest.fixture(scope="session")
def volumes(port, bucky):
    """name -> uint8 volume for every golden `volume` spec (generated once per session)."""
    cache = {}

    def get(spec):
        key = json.dumps(spec, sort_keys=True)
        if key not in cache:
            if spec["kind"] == "bucky":
                cache[key] = bucky
            elif spec["kind"] == "synthA":
                cache[key] = port.synth_a(spec["n"], spec["seed"])
            else:
                cache[key] = port.synth_b(spec["n"], spec["seed"])
        return cache[key]

    return get
