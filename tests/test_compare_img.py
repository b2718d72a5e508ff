"""tools/compare_img: the comparator north_star states the image criterion in (max channel error <= 1/255 on
>= 99.9 % of pixels), CLI of the reference's tools/compare_img.cpp (`compare_img image1 image2`)."""
import os
import struct
import subprocess

import numpy as np

from conftest import build_compare_img, compare_img, write_ppm


def _img(seed=0, w=200, h=100):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def test_identical_and_off_by_one_pass(tmp_path):
    a = _img()
    ok, line = compare_img(a, a, tmp_path)
    assert ok and "identical 100.00000" in line
    b = a.copy()
    b[::2, :, 1] = np.where(b[::2, :, 1] < 255, b[::2, :, 1] + 1, 254)
    ok, _ = compare_img(a, b, tmp_path)
    assert ok


def test_threshold_is_999_permille(tmp_path):
    a = _img(1)
    b = a.copy()
    flat = b.reshape(-1, 3)
    n = len(flat)
    bad = int(n * 0.001)  # exactly 0.1 % wrong: still 99.9 % within tolerance
    flat[:bad, 0] ^= 0x80
    assert compare_img(a, b, tmp_path)[0]
    flat[bad:bad + 5, 2] ^= 0x40
    ok, line = compare_img(a, b, tmp_path)
    assert not ok and "max error" in line
    assert compare_img(a, b, tmp_path, frac=0.99)[0]
    assert compare_img(a, b, tmp_path, tol=255)[0]


def test_formats_and_cli(tmp_path):
    tool = build_compare_img()
    a = _img(2, 37, 21)
    ppm = os.path.join(tmp_path, "a.ppm")
    write_ppm(ppm, a)
    # bottom-up 24-bit TGA and top-down 32-bit TGA of the same pixels
    tga = os.path.join(tmp_path, "a.tga")
    with open(tga, "wb") as f:
        f.write(struct.pack("<BBBHHBHHHHBB", 0, 0, 2, 0, 0, 0, 0, 0, 37, 21, 24, 0))
        f.write(a[::-1].tobytes())
    tga32 = os.path.join(tmp_path, "b.tga")
    with open(tga32, "wb") as f:
        f.write(struct.pack("<BBBHHBHHHHBB", 0, 0, 2, 0, 0, 0, 0, 0, 37, 21, 32, 0x28))
        f.write(np.concatenate([a, np.full((21, 37, 1), 255, np.uint8)], axis=2).tobytes())
    # bottom-up 24-bit BMP (rows padded to 4 bytes)
    bmp = os.path.join(tmp_path, "a.bmp")
    row = (37 * 3 + 3) & ~3
    with open(bmp, "wb") as f:
        f.write(b"BM" + struct.pack("<IHHI", 54 + row * 21, 0, 0, 54))
        f.write(struct.pack("<IiiHHIIiiII", 40, 37, 21, 1, 24, 0, row * 21, 0, 0, 0, 0))
        for y in range(20, -1, -1):
            f.write(a[y].tobytes() + b"\0" * (row - 37 * 3))
    for other in (tga, tga32, bmp):
        r = subprocess.run([tool, ppm, other], capture_output=True, text=True)
        assert r.returncode == 0 and r.stdout.strip().endswith("PASS"), r.stdout
    # the reference tool's own output (mean of every channel per file) is still printed
    r = subprocess.run([tool, ppm, tga], capture_output=True, text=True)
    assert r.stdout.splitlines()[0].startswith(ppm) and "R:" in r.stdout.splitlines()[0]
    # size mismatch fails, usage / unreadable input = 2
    write_ppm(os.path.join(tmp_path, "c.ppm"), _img(3, 36, 21))
    assert subprocess.run([tool, ppm, os.path.join(tmp_path, "c.ppm")], capture_output=True).returncode == 1
    assert subprocess.run([tool, ppm], capture_output=True).returncode == 2
    assert subprocess.run([tool, ppm, os.path.join(tmp_path, "missing.ppm")], capture_output=True).returncode == 2
