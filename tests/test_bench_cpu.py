"""CPU checks of bench.py's host-side pieces: the --tree switch maps to the build flags of include/snail_b200.h, the
committed walker counts behind roofline.algorithmic exist for the headline workload (data, not the oracle), the
reference arm's key set, and that the GPU arm refuses to run without a device instead of falling back."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _bench():
    sys.path.insert(0, ROOT)
    import bench
    return bench


def test_tree_switch_maps_to_the_abi_build_flags(sb):
    b = _bench()
    assert b.tree_flags(sb, "host") == sb.USE_SAH == 4
    assert b.tree_flags(sb, "lbvh") == sb.BUILD_DEVICE == 8
    assert b.tree_flags(sb, "sah") == (sb.BUILD_DEVICE | sb.BUILD_SAH_BINNED) == 24
    hdr = open(os.path.join(ROOT, "include", "snail_b200.h")).read()
    assert "SNL_BUILD_DEVICE" in hdr and "SNL_BUILD_SAH_BINNED" in hdr


def test_headline_walker_counts_are_committed():
    b = _bench()
    wc = b.walker_counts("sponza_like_whitted", 1920, 1080)
    assert wc and wc["rays"] > 7_000_000 and wc["box_tests"] > 50 * wc["rays"] and wc["tri_tests"] > wc["rays"]
    assert b.walker_counts("sponza_like_whitted", 123, 45) is None  # unknown size: the counters build is asked instead


def test_gpu_arm_has_no_cpu_fallback(sb):
    if sb.load_library().snl_device_count() > 0:
        return
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1", "--no-cpu-baseline"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0
    assert not any(line.startswith("{") and "value" in json.loads(line) for line in r.stdout.splitlines() if line.startswith("{"))
