"""CPU tests of the drop-in boundary: libsnail_b200.so loads, exports every symbol include/snail_b200.h declares, and
fails loudly (no CPU fallback) when there is no CUDA device. No compute calls are made here."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def header_symbols():
    text = open(os.path.join(ROOT, "include", "snail_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(snl_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_what_the_binding_lists(sb):
    assert header_symbols() == sorted(sb.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol(sb):
    L = sb.load_library()
    for name in header_symbols():
        assert hasattr(L, name), f"{name} declared in include/snail_b200.h but not exported"
    assert L.snl_abi_version() == 3


def test_counters_build_exports_the_same_abi(sb):
    from snail_b200.build import build_library
    build_library(counters=True)
    L = sb.load_library(counters=True)
    for name in header_symbols():
        assert hasattr(L, name)


def test_struct_layouts_match_the_header(sb):
    assert C.sizeof(sb.Camera) == 13 * 4
    assert C.sizeof(sb.Light) == 36  # src/light.h: 9 floats
    assert C.sizeof(sb.api.Material) == 4 * 11
    assert C.sizeof(sb.Stats) == 10 * 8 + 8  # 10 u64 + float (padded)
    assert sb.NODE_DTYPE.itemsize == 32 and sb.TRI_DTYPE.itemsize == 64 and sb.SHTRI_DTYPE.itemsize == 64


def test_no_cpu_fallback_without_a_device(sb):
    L = sb.load_library()
    if L.snl_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(sb.SnlError) as e:
        sb.Scene(0)
    assert "no CUDA device" in str(e.value)


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under snail_b200/ may import, link or execute oracle/."""
    pkg = os.path.join(ROOT, "snail_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "snailref" not in src and "oracle/" not in src.replace("never touches oracle/", ""), f


def test_header_is_plain_c_and_the_c_caller_links(sb, tmp_path):
    """include/snail_b200.h must be usable from C (no C++-isms, no torch / CUDA types), and the plain-C multi-GPU caller
    of tests/c links against the library and refuses to run - cleanly, with a message - on a box without a device."""
    import subprocess
    inc, lib_dir = os.path.join(ROOT, "include"), os.path.join(ROOT, "snail_b200")
    probe = tmp_path / "probe.c"
    probe.write_text('#include "snail_b200.h"\nint main(void) { return sizeof(snl_camera) + sizeof(snl_stats) > 0 ? 0 : 1; }\n')
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-fsyntax-only", "-I", inc, str(probe)], check=True)
    exe = str(tmp_path / "render_multi_test")
    subprocess.run(["gcc", "-std=c11", "-O1", "-Wall", "-Werror", "-I", inc, os.path.join(ROOT, "tests", "c", "render_multi_test.c"),
                    "-L", lib_dir, "-lsnail_b200", "-lm", f"-Wl,-rpath,{lib_dir}", "-o", exe], check=True)
    if sb.load_library().snl_device_count() > 0:
        return  # the GPU suite runs it for real (tests/test_gpu_multi.py)
    r = subprocess.run([exe, os.path.join(ROOT, "data", "box.obj"), "1", "64", "64", "1"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "0 visible" in r.stderr, (r.returncode, r.stdout, r.stderr)
