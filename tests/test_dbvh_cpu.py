"""Host side of the instanced two-level BVH (snail_b200/csrc/snl_host_dbvh.cpp) against the committed goldens, which
hold what the reference's ObjectInstance::ComputeBBox + DBVH::Construct produced for the same instances
(tools/make_golden.py instanced_cases): node records bit for bit (sha256), the order std::partition leaves the
instances in, and their boxes. No device needed."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

MANIFEST = json.load(open(os.path.join(GOLDEN, "manifest.json")))


def find_scene(name):
    for d in (os.path.join(ROOT, "assets"), os.path.join(ROOT, "oracle", "_ref")):
        p = os.path.join(d, "scenes", name)
        if os.path.exists(p):
            return p
    pytest.skip(f"scenes/{name}: scene assets not present")


@pytest.mark.parametrize("name", list(MANIFEST.get("instanced", {})))
def test_host_dbvh_is_the_reference_dbvh(sb, name):
    c = MANIFEST["instanced"][name]
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    hs = sb.HostScene(find_scene(c["obj"]))
    bmin, bmax = hs.bbox()
    nodes, elems, boxes = sb.host_dbvh_build(g["instances"], bmin, bmax)
    assert len(nodes) == c["n_nodes"]
    assert (elems == g["dbvh_elements"]).all()  # same permutation of the instances
    assert (boxes == g["dbvh_boxes"]).all()     # ComputeBBox bit for bit
    assert hashlib.sha256(np.ascontiguousarray(nodes).tobytes()).hexdigest() == c["sha256_dbvh_nodes"]
