"""Generates tests/golden/*.npz from the COMPILED REFERENCE (oracle/_ref, built
from the headers under /root/reference by oracle/Makefile). Run in the build
container only (needs /root/reference for the real fixture frame):

    make -C oracle && python tests/golden/make_golden.py

Each file holds, for one input frame and intrinsic: the u16 depth input, and
for input types pinhole-depth (kind 0) / pinhole-range (kind 1) the float range
image, every float feature image and its 16-bit quantisation exactly as the
reference computes them (Real = float, CLI path)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import (FEATURE_NAMES, KINECT, Oracle, RefLib,  # noqa: E402
                        kinect_scaled)


def f32k(k):
    return {**k, **{n: float(np.float32(k[n])) for n in ("fx", "fy", "cx", "cy")}}


def case(ref, depth, k, kinds=(0, 1)):
    out = {"depth": depth,
           "intrinsic": np.array([k["w"], k["h"], k["fx"], k["fy"], k["cx"], k["cy"]],
                                 np.float64)}
    for kind in kinds:
        rng = ref.laserscan(depth, k) if kind == 0 else depth.astype(np.float32)
        out[f"k{kind}_range"] = rng
        out[f"k{kind}_range_u16"] = ref.quantise(0, rng, rng)
        for f in range(1, 9):
            feat = ref.feature(f, rng, k)
            out[f"k{kind}_{FEATURE_NAMES[f]}"] = feat
            out[f"k{kind}_{FEATURE_NAMES[f]}_u16"] = ref.quantise(f, feat, rng, -20.0, 20.0)
    return out


def main():
    ref, orc = RefLib(), Oracle()
    print(ref.info())
    # 1. synthetic 96x64, Kinect intrinsic scaled to the frame
    k = kinect_scaled(96, 64)
    d = orc.synth_depth(0xD2C00000, 0, 96, 64)
    np.savez_compressed(os.path.join(HERE, "synth_96x64.npz"), **case(ref, d, k))
    # 2. odd width (scalar path), small
    k = kinect_scaled(61, 37)
    d = orc.synth_depth(0xD2C00000, 1, 61, 37)
    np.savez_compressed(os.path.join(HERE, "synth_61x37.npz"), **case(ref, d, k))
    # 3. a 160x120 crop of the reference's own Kinect fixture frame
    #    (src/test/depth2x/data0-depth.png), with an intrinsic of that size
    import cv2
    img = cv2.imread("/root/reference/src/test/depth2x/data0-depth.png", cv2.IMREAD_UNCHANGED)
    crop = np.ascontiguousarray(img[200:320, 400:560])
    k = f32k(dict(w=160, h=120, fx=519.226, fy=522.23, cx=79.462, cy=62.737))
    np.savez_compressed(os.path.join(HERE, "kinect_crop_160x120.npz"), **case(ref, crop, k, kinds=(0,)))
    # 4. known-answer scalars: phi / pixel_to_sphere / bearing_angle at sample points
    k = f32k(KINECT)
    pts = [(0, 0), (1, 1), (479, 272), (480, 273), (959, 539), (100, 500), (900, 20)]
    rays = np.array([ref.pixel_to_sphere(k, u, v) for u, v in pts], np.float32)
    pairs = [(1, 1, 0, 1), (1, 1, 1, 0), (1, 1, 0, 0), (1, 1, 0, 2), (480, 270, 479, 270),
             (959, 539, 958, 538), (10, 10, 12, 10), (10, 9, 10, 11), (9, 9, 11, 11)]
    phis = np.array([ref.phi(k, *p) for p in pairs], np.float32)
    ba_in = [(1.0, 1.0, 0.5), (5.0, 5.0, 0.70710677), (2.4, 6.7, 0.7193398),
             (1000.0, 1001.0, 0.99999881), (1500.5, 1499.25, 0.99999797),
             (800.0, 2400.0, 0.99999881), (1.0, 1.4142135, 0.70710677)]
    ba = np.array([ref.bearing_angle(*x) for x in ba_in], np.float32)
    np.savez_compressed(os.path.join(HERE, "kat_scalars.npz"),
                        pts=np.array(pts), rays=rays, pairs=np.array(pairs), phis=phis,
                        ba_in=np.array(ba_in, np.float32), ba=ba)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
