"""CPU: the depth2x command line, the PNG codec and the intrinsic parser of the
C++ host layer. The argument cases are those of the reference's shell test
src/test/depth2x/test_depth2x.sh:22-65 that need no conversion; the conversions
themselves are in test_depth2x_gpu.py."""
import os
import subprocess

import cv2
import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(os.path.dirname(HERE), "depth-conversions_b200", "bin")
DEPTH2X = os.path.join(BIN, "depth2x")
SELFTEST = os.path.join(BIN, "host_selftest")
KINECT = "960 540\n519.226 0.000000 479.462\n0.000000 522.23 272.737\n0.000000 0.000000 1.000000\n"


def run(args, cwd):
    return subprocess.run([DEPTH2X] + args, cwd=cwd, capture_output=True, text=True, timeout=120)


@pytest.fixture()
def work(tmp_path):
    (tmp_path / "kinect_intrinsic.txt").write_text(KINECT)
    cv2.imwrite(str(tmp_path / "data0-depth.png"), np.full((540, 960), 1000, np.uint16))
    return str(tmp_path)


def test_tools_are_built():
    assert os.access(DEPTH2X, os.X_OK) and os.access(SELFTEST, os.X_OK)


def test_argument_validation(work):
    # test_depth2x.sh:22-25 no arguments
    assert run([], work).returncode != 0
    # :27-30 --version works and prints "<argv0> v1.0.0" (version_printer.cpp:9-12)
    r = run(["--version"], work)
    assert r.returncode == 0 and r.stdout.strip().endswith("depth2x v1.0.0")
    # :32-35 a sub-command is required
    assert run(["-c", "kinect_intrinsic.txt", "-i", "data0-depth.png", "-s", "0", "-e", "4"],
               work).returncode != 0
    # :37-40 intrinsic missing for a sub-command that needs it
    r = run(["bearing", "-i", "data0-depth.png", "-s", "0", "-e", "4"], work)
    assert r.returncode == 1 and "Could not load intrinsic calibration" in r.stderr
    # :44-53 bearing without any output pattern, sub-command last or first
    for args in (["-c", "kinect_intrinsic.txt", "-i", "data0-depth.png", "-s", "0", "-e", "4", "bearing"],
                 ["bearing", "-c", "kinect_intrinsic.txt", "-i", "data0-depth.png", "-s", "0", "-e", "4"]):
        r = run(args, work)
        assert r.returncode == 1
        assert "Severe problem occured while system-setup." in r.stderr
        assert "Missing output pattern for at least one bearing direction" in r.stderr
    # :56-59 unsupported camera model
    assert run(["flexion", "-m", "scaramuzza", "-c", "kinect_intrinsic.txt", "-i", "data0-depth.png",
                "-s", "0", "-e", "4", "-o", "x{}.png"], work).returncode != 0
    # required flags, unknown flags, a calibration file that does not exist, bad --type
    assert run(["flexion", "-c", "kinect_intrinsic.txt", "-i", "d{}.png", "-s", "0", "-o", "x"],
               work).returncode != 0
    assert run(["flexion", "-c", "kinect_intrinsic.txt", "-i", "d{}.png", "-s", "0", "-e", "1"],
               work).returncode != 0
    assert run(["flexion", "--frobnicate", "-i", "d{}.png", "-s", "0", "-e", "1", "-o", "x"],
               work).returncode != 0
    assert run(["flexion", "-c", "nope.txt", "-i", "d{}.png", "-s", "0", "-e", "1", "-o", "x"],
               work).returncode != 0
    assert run(["flexion", "-t", "fisheye", "-c", "kinect_intrinsic.txt", "-i", "d{}.png", "-s", "0",
                "-e", "1", "-o", "x"], work).returncode != 0
    assert run(["--help"], work).returncode == 0


@pytest.mark.parametrize("text", [
    "960 540 7\n519.226 0 479.462\n0 522.23 272.737\n",      # trailing token on line 1
    "960 540\n519.226 0 479.462 1\n0 522.23 272.737\n",      # trailing token on line 2
    "960 540\n-519.226 0 479.462\n0 522.23 272.737\n",       # fx <= 0
    "960\n519.226 0 479.462\n0 522.23 272.737\n",            # h missing
    "960 540\n519.226 0 479.462\n0 522.23\n",                # cy missing
    "",
])
def test_intrinsic_parser_is_strict(work, text):
    """io/intrinsics.h:48-98: each line must be consumed exactly."""
    with open(os.path.join(work, "bad.txt"), "w") as f:
        f.write(text)
    r = run(["flexion", "-c", "bad.txt", "-i", "data{}-depth.png", "-s", "0", "-e", "0", "-o", "o{}.png"],
            work)
    assert r.returncode == 1 and "Could not load intrinsic calibration" in r.stderr


def test_png_codec_roundtrip(tmp_path):
    """This build's PNG reader/writer against OpenCV's (16 bit grey, all five
    filter types on the read side, odd sizes)."""
    rng = np.random.default_rng(7)
    cases = {
        "noise": rng.integers(0, 65536, (37, 61), dtype=np.uint16),
        "ramp": (np.arange(540 * 960, dtype=np.uint32).reshape(540, 960) % 65536).astype(np.uint16),
        "one": np.array([[65535]], np.uint16),
        "zeros": np.zeros((5, 1), np.uint16),
        # the writer's own deflate encoder (io/fast_deflate.cpp): several 256 KB blocks,
        # runs that cross block boundaries, incompressible blocks (stored), run lengths
        # around the 258 limit, a constant image
        "big_zeros": np.zeros((540, 960), np.uint16),
        "big_noise": rng.integers(0, 65536, (540, 960), dtype=np.uint16),
        "constant": np.full((300, 500), 0x1234, np.uint16),
        "holes": np.where(rng.random((540, 960)) < 0.3, 0,
                          rng.integers(700, 2600, (540, 960))).astype(np.uint16),
        "runs": np.repeat(rng.integers(0, 65536, (64, 40), dtype=np.uint16),
                          rng.integers(1, 300, 40), axis=1),
        "blocky": np.kron(rng.integers(0, 3, (34, 60)), np.ones((16, 16), np.int64)).astype(np.uint16) * 999,
    }
    for name, img in cases.items():
        a, b = str(tmp_path / f"{name}_a.png"), str(tmp_path / f"{name}_b.png")
        cv2.imwrite(a, img)
        assert subprocess.run([SELFTEST, "png", a, b]).returncode == 0
        back = cv2.imread(b, cv2.IMREAD_UNCHANGED)
        assert back.dtype == np.uint16 and back.shape == img.shape and np.array_equal(back, img)
    # adaptive filtering on the writer side (PIL picks filters per row)
    from PIL import Image
    p = str(tmp_path / "pil.png")
    Image.fromarray(cases["ramp"]).save(p, optimize=True)
    assert subprocess.run([SELFTEST, "png", p, p + ".out.png"]).returncode == 0
    assert np.array_equal(cv2.imread(p + ".out.png", cv2.IMREAD_UNCHANGED), cases["ramp"])
    # 8 bit and colour images are not depth images (io/image.h:28-29): rejected
    cv2.imwrite(str(tmp_path / "u8.png"), np.zeros((4, 4), np.uint8))
    assert subprocess.run([SELFTEST, "png", str(tmp_path / "u8.png"), "/dev/null"]).returncode != 0
    cv2.imwrite(str(tmp_path / "rgb.png"), np.zeros((4, 4, 3), np.uint16))
    assert subprocess.run([SELFTEST, "png", str(tmp_path / "rgb.png"), "/dev/null"]).returncode != 0
    assert subprocess.run([SELFTEST, "png", str(tmp_path / "missing.png"), "/dev/null"]).returncode != 0


def test_no_cpu_fallback(work):
    """Without a B200 a conversion must fail loudly, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by test_depth2x_gpu.py")
    r = run(["flexion", "-c", "kinect_intrinsic.txt", "-i", "data{}-depth.png", "-s", "0", "-e", "0",
             "-o", "out{}.png"], work)
    assert r.returncode == 1
    assert "no CUDA device" in r.stderr or "CUDA" in r.stderr
    assert "Could not process index \"0\"" in r.stderr
    assert not os.path.exists(os.path.join(work, "out0.png"))


def test_png_writer_deflate_fuzz(tmp_path):
    """io/fast_deflate.cpp (run-length + dynamic Huffman) through the PNG writer,
    decoded by OpenCV (libpng + zlib inflate): random sizes and run structures."""
    rs = np.random.RandomState(1)
    a, b = str(tmp_path / "a.png"), str(tmp_path / "b.png")
    for t in range(24):
        h, w = rs.randint(1, 200), rs.randint(1, 700)
        kind = t % 6
        if kind == 0:
            img = rs.randint(0, 65536, (h, w))
        elif kind == 1:
            img = np.zeros((h, w))
        elif kind == 2:
            img = np.repeat(rs.randint(0, 65536, (h, 1)), w, axis=1)
        elif kind == 3:
            img = np.where(rs.rand(h, w) < 0.5, 0, rs.randint(0, 4, (h, w)) * 257)
        elif kind == 4:
            img = np.cumsum(rs.randint(0, 2, (h, w)), axis=1) * 3
        else:
            img = np.kron(rs.randint(0, 65536, (h // 8 + 1, w // 8 + 1)), np.ones((8, 8), np.int64))[:h, :w]
        img = img.astype(np.uint16)
        cv2.imwrite(a, img)
        assert subprocess.run([SELFTEST, "png", a, b]).returncode == 0
        back = cv2.imread(b, cv2.IMREAD_UNCHANGED)
        assert back is not None and back.shape == img.shape and np.array_equal(back, img), (t, kind, h, w)
    # the zlib path stays selectable
    env = dict(os.environ, DC_PNG_ZLIB="1")
    assert subprocess.run([SELFTEST, "png", a, b], env=env).returncode == 0
    assert np.array_equal(cv2.imread(b, cv2.IMREAD_UNCHANGED), img)


def test_png_reader_inflate_fuzz():
    """io/fast_inflate.cpp (the PNG reader's table-driven inflate) against streams zlib
    writes at every level / strategy / window / flush pattern, from all-literal to
    all-match data; damaged streams are refused or accepted exactly as zlib does."""
    for seed, iters, env in ((1, 500, None), (2, 500, dict(os.environ, DC_INFLATE_SCALAR="1"))):
        r = subprocess.run([SELFTEST, "inflatefuzz", str(seed), str(iters)], capture_output=True, text=True,
                           env=env)
        assert r.returncode == 0 and r.stdout.startswith("ok %d " % iters), r.stdout + r.stderr


def test_png_reader_other_writers(tmp_path):
    """The reader on files from other encoders: libpng at every compression level and
    zlib strategy (many 8 KB IDAT chunks, all five row filters), PIL (one big IDAT), and a
    truncated file (refused by both the fast path and the zlib fallback)."""
    rng = np.random.default_rng(3)
    y, x = np.mgrid[0:270, 0:480]
    img = (3000 + 800 * np.sin(x / 37.0) + 500 * np.cos(y / 23.0) + rng.normal(0, 5, x.shape)).astype(np.uint16)
    img[rng.random(img.shape) < 0.05] = 0
    img[100:140, 200:330] = 0
    a, b = str(tmp_path / "a.png"), str(tmp_path / "b.png")
    strategies = [cv2.IMWRITE_PNG_STRATEGY_DEFAULT, cv2.IMWRITE_PNG_STRATEGY_FILTERED,
                  cv2.IMWRITE_PNG_STRATEGY_HUFFMAN_ONLY, cv2.IMWRITE_PNG_STRATEGY_RLE,
                  cv2.IMWRITE_PNG_STRATEGY_FIXED]
    for level in range(10):
        for strategy in strategies:
            assert cv2.imwrite(a, img, [cv2.IMWRITE_PNG_COMPRESSION, level, cv2.IMWRITE_PNG_STRATEGY, strategy])
            for env, which in ((None, "fast"), (dict(os.environ, DC_PNG_ZLIB="1"), "zlib")):
                r = subprocess.run([SELFTEST, "png", a, b], env=env, capture_output=True, text=True)
                # (the fast path must take these itself, not lean on its fallback)
                assert r.returncode == 0 and r.stdout.strip() == "inflate " + which, (level, strategy, r.stdout)
                assert np.array_equal(cv2.imread(b, cv2.IMREAD_UNCHANGED), img), (level, strategy)
    from PIL import Image
    Image.fromarray(img).save(a, compress_level=9)
    assert subprocess.run([SELFTEST, "png", a, b]).returncode == 0
    assert np.array_equal(cv2.imread(b, cv2.IMREAD_UNCHANGED), img)
    data = open(a, "rb").read()
    open(a, "wb").write(data[: len(data) // 2])
    assert subprocess.run([SELFTEST, "png", a, b]).returncode != 0
