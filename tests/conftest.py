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


def pytest_collection_modifyitems(config, items):
    """A stuck test (a lost collective, a wedged device) must fail, not hold the box: 20 minutes per test where
    pytest-timeout is installed (the slowest test takes about one)."""
    if config.pluginmanager.hasplugin("timeout"):
        for item in items:
            if item.get_closest_marker("timeout") is None:
                item.add_marker(pytest.mark.timeout(1200))


COMPARE_IMG = os.path.join(ROOT, "tools", "compare_img")


def build_compare_img():
    """tools/compare_img (north_star's comparator, CLI of the reference's tools/compare_img.cpp) built in place."""
    import subprocess
    src = COMPARE_IMG + ".cpp"
    if not os.path.exists(COMPARE_IMG) or os.path.getmtime(COMPARE_IMG) < os.path.getmtime(src):
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", COMPARE_IMG, src], check=True)
    return COMPARE_IMG


def write_ppm(path, img):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(img.tobytes())


def compare_img(a, b, tmp_path, tol=1, frac=0.999):
    """Runs tools/compare_img on two (H, W, 3) uint8 frames; returns (passed, the tool's output)."""
    import subprocess
    tool = build_compare_img()
    pa, pb = os.path.join(str(tmp_path), "a.ppm"), os.path.join(str(tmp_path), "b.ppm")
    write_ppm(pa, a)
    write_ppm(pb, b)
    r = subprocess.run([tool, pa, pb, "--tol", str(tol), "--frac", str(frac)], capture_output=True, text=True)
    assert r.returncode in (0, 1), r.stdout + r.stderr
    return r.returncode == 0, r.stdout.strip().splitlines()[-2] if r.stdout.count("\n") >= 2 else r.stdout


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
