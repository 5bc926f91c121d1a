"""Debug aid: which pixels of a golden case differ, per feature."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from __graft_entry__ import load_pkg
P = load_pkg()
z = np.load(os.path.join(ROOT, "tests", "golden", sys.argv[1]))
w, h, fx, fy, cx, cy = z["intrinsic"]
with P.Context(0) as ctx:
    ctx.set_pinhole(int(w), int(h), float(fx), float(fy), float(cx), float(cy))
    d = z["depth"]
    g, m = ctx.convert(P.DC_CURVATURE, d, P.DC_IN_U16_DEPTH, P.DC_OUT_F32, want=(True, True))[:2]
    r = z["k0_range"]
    for name, got in (("gauss", g), ("mean", m)):
        want = z["k0_" + name]
        bad = np.argwhere(got.view(np.uint32) != want.view(np.uint32))
        print(name, len(bad), "differ")
        for v, u in bad[:5]:
            print("  at", v, u, repr(got[v, u]), repr(want[v, u]), hex(got.view(np.uint32)[v, u]), hex(want.view(np.uint32)[v, u]))
            print(r[v - 1:v + 2, u - 1:u + 2])
