"""Builds snail_b200/libsnail_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

`python -m snail_b200.build [--counters] [--force]`
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsnail_b200.so")
LIB_COUNTERS = os.path.join(HERE, "libsnail_b200_counters.so")

SOURCES = ["snl_kernels.cu", "snl_wavefront.cu", "snl_build.cu", "snl_build_sah.cu", "snl_capi.cu", "snl_comm.cu", "snl_peaks.cu", "snl_host_scene.cpp", "snl_host_dbvh.cpp"]
HEADERS = ["snl_device.cuh", "snl_kernels.h", "snl_scene.h", "snl_scene_dev.h", "snl_types.h", "snl_host_scene.h",
           os.path.join("..", "..", "include", "snail_b200.h")]

# -fmad=false / -ffp-contract=off: every float expression keeps the reference's operation order and rounding,
# which is what makes hit records bit-identical to the reference built on veclib's scalar back-end.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "--use_fast_math=false"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, counters: bool = False, verbose: bool = False, extra=(), suffix: str = "") -> str:
    """extra / suffix: experimental variants (extra nvcc flags -> libsnail_b200<suffix>.so, picked up through SNL_LIB)."""
    target = LIB_COUNTERS if counters else LIB
    if suffix:
        target = target[:-3] + suffix + ".so"
    if not force and not _stale(target):
        return target
    flags = [f for f in NVCC_FLAGS if f != "--use_fast_math=false"]
    if "-fmad=true" in extra:  # A/B only: contracts a*b+c everywhere, t / u / v are no longer bit-identical with the reference
        flags = [f for f in flags if f != "-fmad=false"]
    if counters:
        flags += ["-DSNL_COUNTERS"]
    flags += list(extra)
    if verbose:
        flags += ["-Xptxas", "-v"]
    # one nvcc per translation unit, in parallel; objects are cached by (flags, source + header contents)
    import hashlib
    from concurrent.futures import ThreadPoolExecutor
    inc = ["-I", os.path.join(HERE, "..", "include")]
    hdr_digest = hashlib.sha256(b"".join(open(os.path.join(CSRC, h), "rb").read() for h in HEADERS)).hexdigest()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        key = hashlib.sha256((" ".join(flags) + hdr_digest).encode() + open(os.path.join(CSRC, src), "rb").read()).hexdigest()[:20]
        obj = os.path.join(objdir, f"{os.path.splitext(src)[0]}.{key}.o")
        if os.path.exists(obj) and not verbose:
            return obj, ""
        cmd = [_nvcc(), "-c", *flags, *inc, os.path.join(CSRC, src), "-o", obj + ".tmp"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
        os.replace(obj + ".tmp", obj)
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        results = list(ex.map(compile_one, SOURCES))
    if verbose:
        print("".join(r[1] for r in results))
    keep = {os.path.basename(r[0]) for r in results}
    for f in os.listdir(objdir):  # drop objects of older source versions of the same configuration
        stem = f.split(".")[0]
        if f.endswith(".o") and f not in keep and any(stem == os.path.splitext(x)[0] for x in SOURCES) and \
                os.path.getmtime(os.path.join(objdir, f)) < __import__("time").time() - 6 * 3600:
            os.remove(os.path.join(objdir, f))
    cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", *[r[0] for r in results], "-o", target + ".tmp", "-lz"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    os.replace(target + ".tmp", target)
    return target


if __name__ == "__main__":
    extra = [a for a in sys.argv[1:] if a.startswith("-D") or a.startswith("-maxrregcount") or a == "-fmad=true"]
    suffix = next((a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--suffix=")), "")
    path = build_library(force="--force" in sys.argv, counters="--counters" in sys.argv, verbose="-v" in sys.argv,
                         extra=extra, suffix=suffix)
    print(path)
