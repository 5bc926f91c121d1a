#!/usr/bin/env python
"""How far may the bearing ratio deviate from the reference's exactly rounded one?
Runs bearing x4 of the library named by DC_LIB (a variant built with
-DDC_APPROX_RATIO=1, or the product library) on the real Kinect crop of
tests/golden and on synthetic 960x540 frames, and compares with the golden vectors /
the oracle: pixels outside the contract (1e-5 relative for float, 1 LSB for u16)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import importlib.util
spec = importlib.util.spec_from_file_location("pkg", os.path.join(ROOT, "depth-conversions_b200", "__init__.py"))
pkg = importlib.util.module_from_spec(spec); spec.loader.exec_module(pkg)
from oracle_lib import Oracle, kinect_scaled

def stats(name, got32, want32, got16, want16):
    nz = want32 != 0
    rel = np.abs(got32.astype(np.float64) - want32) / np.maximum(np.abs(want32), 1e-30)
    rel[~nz] = 0
    d16 = np.abs(got16.astype(np.int32) - want16.astype(np.int32))
    print(f"{name}: pixels {nz.sum()}  bit-equal {(got32.view(np.uint32) == want32.view(np.uint32))[nz].mean():.4f}  "
          f"rel>1e-5 {(rel > 1e-5).sum()}  rel>1e-4 {(rel > 1e-4).sum()}  max_rel {rel.max():.3e}  "
          f"u16 off>1 {(d16 > 1).sum()}  max_u16 {d16.max()}  zero-mask equal {np.array_equal(got32 == 0, want32 == 0)}")

ctx = pkg.Context(0)
orc = Oracle()
z = np.load(os.path.join(ROOT, "tests", "golden", "kinect_crop_160x120.npz"))
w, h, fx, fy, cx, cy = z["intrinsic"]
ctx.set_pinhole(int(w), int(h), float(fx), float(fy), float(cx), float(cy))
f32 = ctx.convert(pkg.DC_BEARING, z["depth"], pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32, want=(True,) * 4)
u16 = ctx.convert(pkg.DC_BEARING, z["depth"], pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
print("library", pkg.LIB_PATH, "kernel", ctx.last_kernel)
for d, n in enumerate("hvda"):
    stats(f"kinect_crop_160x120 bearing_{n}", f32[d], z[f"k0_bearing_{n}"], u16[d], z[f"k0_bearing_{n}_u16"])
W, H = 960, 540
k = kinect_scaled(W, H)
ctx.set_pinhole(W, H, k["fx"], k["fy"], k["cx"], k["cy"])
frames = np.stack([orc.synth_depth(0xD2C00002, f, W, H) for f in range(2)])
# a depth discontinuity (foreground object) as real scenes have them
frames[1, 100:300, 200:500] = np.where(frames[1, 100:300, 200:500] > 0, frames[1, 100:300, 200:500] // 2, 0)
f32 = ctx.convert(pkg.DC_BEARING, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32, want=(True,) * 4)
u16 = ctx.convert(pkg.DC_BEARING, frames, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_U16, want=(True,) * 4)
for i in range(2):
    rng = orc.laserscan(frames[i], k)
    for d, n in enumerate("hvda"):
        want = orc.feature(1 + d, rng, k)
        stats(f"synth_960x540 frame {i}{' (with a foreground box)' if i else ''} bearing_{n}", f32[d][i], want, u16[d][i], orc.quantise(1 + d, want, rng))
