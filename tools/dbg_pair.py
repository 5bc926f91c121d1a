"""Debug aid: pair kernels (u16 input) against the one-frame kernels (f32 range input)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_pkg
from oracle_lib import Oracle, kinect_scaled
P = load_pkg(); orc = Oracle()
w, h = int(sys.argv[1]) if len(sys.argv) > 1 else 160, int(sys.argv[2]) if len(sys.argv) > 2 else 120
n = int(sys.argv[3]) if len(sys.argv) > 3 else 3
k = kinect_scaled(w, h)
frames = np.stack([orc.synth_depth(0xD2C00001, f, w, h) for f in range(n)])
with P.Context(0) as ctx:
    ctx.set_pinhole(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
    rng = ctx.convert(P.DC_RANGE, frames, P.DC_IN_U16_DEPTH, P.DC_OUT_F32)[0]
    for feat, name, want in ((P.DC_BEARING, "bearing", (True,) * 4), (P.DC_CURVATURE, "curv", (True, True)),
                             (P.DC_MAX_CURVE, "maxcurve", (True, True)), (P.DC_FLEXION, "flexion", (True,))):
        for ok, okn in ((P.DC_OUT_F32, "f32"), (P.DC_OUT_U16, "u16")):
            a = ctx.convert(feat, frames, P.DC_IN_U16_DEPTH, ok, want=want)
            b = ctx.convert(feat, rng, P.DC_IN_F32_RANGE, ok, want=want)
            for i, (x, y) in enumerate(zip(a, b)):
                if x is None: continue
                if ok == P.DC_OUT_F32:
                    bad = x.view(np.uint32) != y.view(np.uint32)
                else:
                    bad = x != y
                msg = f"{name}[{i}] {okn}: {int(bad.sum())} of {bad.size} differ"
                if bad.any():
                    idx = np.argwhere(bad)
                    fr, vv, uu = idx[:, 0], idx[:, 1], idx[:, 2]
                    msg += f" frames {np.unique(fr)[:8]} rows {np.unique(vv)[:12]} cols {np.unique(uu)[:12]}"
                    f0, v0, u0 = idx[0]
                    msg += f" first ({f0},{v0},{u0}): {x[f0, v0, u0]} vs {y[f0, v0, u0]}"
                    if ok == P.DC_OUT_F32:
                        rel = np.abs(x.astype(np.float64) - y) / np.maximum(np.abs(y), 1e-30)
                        msg += f" max rel {rel[bad].max():.3e}"
                        big = np.argwhere(bad & (rel > 3e-6))
                        if len(big):
                            msg += f"\n    big: n={len(big)} col%4 hist {np.bincount(big[:, 2] % 4, minlength=4)} row%18 hist {np.bincount(big[:, 1] % 18, minlength=18)} frame hist {np.bincount(big[:, 0])}"
                            f0, v0, u0 = big[0]
                            msg += f" first big ({f0},{v0},{u0}): {x[f0, v0, u0]!r} vs {y[f0, v0, u0]!r}"
                    else:
                        msg += f" max |d| {np.abs(x.astype(int) - y.astype(int)).max()}"
                print(msg)
