import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

REF_DIR = os.path.join(ROOT, "oracle", "_ref")
SCENES = os.path.join(REF_DIR, "scenes")
TEXDIR = os.path.join(REF_DIR, "data")
GOLDEN = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(ROOT, "data")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def scene_path(name):
    p = os.path.join(SCENES, name + ".obj")
    if not os.path.exists(p):
        pytest.skip(f"{p} missing (oracle/_ref assets are copied by oracle/build_ref.sh)")
    return p


def need_oracle(variant="scalar"):
    import snailref
    if not snailref.available(variant):
        pytest.skip(f"oracle/_ref/libsnailref_{variant}.so not built")
    return snailref


@pytest.fixture(scope="session")
def sb():
    from snail_b200.build import build_library
    build_library()
    import snail_b200
    return snail_b200
