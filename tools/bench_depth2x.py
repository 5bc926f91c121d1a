#!/usr/bin/env python
"""End-to-end (PNG in -> PNG out) timing of bin/depth2x against the reference's
CPU path on the same files: SURVEY.md section 8d, config 1 (flexion, 960x540,
100 frames) and, scaled down, config 5 (all features, 1920x1080).

The reference's own depth2x cannot be built here (needs OpenCV C++); its CPU arm
is therefore the compiled reference headers (oracle/_ref) driven frame-parallel
from a thread pool with cv2 as the PNG codec (cv2.imread -> depth_to_laserscan +
depth_to_X + convert_X -> cv2.imwrite; cv2 and ctypes release the GIL), i.e. the
work of batch_converter::process_index (src/apps/util/batch_converter.cpp:12-28)
per index. Wall clock on both arms, files on the same tmpfs/disk.

Prints one JSON line per (feature, arm)."""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import (F_BEARING_H, F_FLEXION, F_GAUSS, F_MAX_CURVE, F_MEAN, F_RANGE, Oracle,  # noqa: E402
                        RefLib, kinect_scaled, ref_available)

DEPTH2X = os.path.join(ROOT, "depth-conversions_b200", "bin", "depth2x")
SUBCMD = {
    "flexion": (["flexion", "-o", "out/flexion-{}.png"], [F_FLEXION]),
    "bearing": (["bearing", "--horizontal", "out/bh-{}.png", "--vertical", "out/bv-{}.png",
                 "--diagonal", "out/bd-{}.png", "--anti-diagonal", "out/ba-{}.png"],
                [F_BEARING_H, F_BEARING_H + 1, F_BEARING_H + 2, F_BEARING_H + 3]),
    "range": (["range", "-o", "out/range-{}.png"], [F_RANGE]),
    "max-curve": (["max-curve", "-o", "out/mc-{}.png"], [F_MAX_CURVE]),
    "gauss-curvature": (["gauss-curvature", "-o", "out/g-{}.png"], [F_GAUSS]),
    "mean-curvature": (["mean-curvature", "-o", "out/m-{}.png"], [F_MEAN]),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--size", default="960x540")
    ap.add_argument("--features", default="flexion")
    ap.add_argument("--devices", type=int, default=1)
    ap.add_argument("--cpu-frames", type=int, default=0, help="frames of the CPU arm (0: same)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--workdir", default="")
    a = ap.parse_args()
    w, h = (int(x) for x in a.size.split("x"))
    k = kinect_scaled(w, h)
    orc = Oracle()
    work = a.workdir or tempfile.mkdtemp(prefix="depth2x_bench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    os.makedirs(os.path.join(work, "out"), exist_ok=True)
    with open(os.path.join(work, "intrinsic.txt"), "w") as f:
        f.write(f"{w} {h}\n{k['fx']!r} 0.0 {k['cx']!r}\n0.0 {k['fy']!r} {k['cy']!r}\n0.0 0.0 1.0\n")
    distinct = [orc.synth_depth(0xD2C00001, i, w, h) for i in range(8)]
    for i in range(a.frames):
        cv2.imwrite(os.path.join(work, f"data{i}-depth.png"), distinct[i % 8])
    in_bytes = sum(os.path.getsize(os.path.join(work, f"data{i}-depth.png")) for i in range(a.frames))
    ncpu = os.cpu_count()
    cpu_model = next((l.split(":")[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")), "?")
    for feat in a.features.split(","):
        sub, fids = SUBCMD[feat]
        # ---- GPU arm: bin/depth2x ----
        cmd = [DEPTH2X, "-c", "intrinsic.txt", "-i", "data{}-depth.png", "-s", "0", "-e",
               str(a.frames - 1), "--devices", str(a.devices)] + sub
        best = None
        for rep in range(3):  # first run pages the library in; report the best of 3
            t0 = time.perf_counter()
            r = subprocess.run(cmd, cwd=work, capture_output=True, text=True)
            dt = time.perf_counter() - t0
            assert r.returncode == 0, r.stdout + r.stderr
            best = dt if best is None else min(best, dt)
        inner = [l.strip() for l in (r.stdout + r.stderr).splitlines() if "took" in l]
        # the pipeline alone (reader/GPU/writer stages, without CUDA start-up, which varies
        # between 0.2 s and 4 s on a fresh box): DEPTH2X_PROFILE timestamps
        rp = subprocess.run(cmd, cwd=work, capture_output=True, text=True,
                            env=dict(os.environ, DEPTH2X_PROFILE="1"))
        marks = {}
        for l in rp.stderr.splitlines():
            if l.startswith("[profile"):
                t = float(l[len("[profile"):l.index("s]")])
                marks[l[l.index("]") + 2:].rsplit(" ", 1)[0]] = t
        pipe = marks.get("pipeline drained", 0.0) - marks.get("dc_create end", 0.0)
        print(json.dumps({"arm": "depth2x (B200)", "feature": feat, "size": a.size, "frames": a.frames,
                          "devices": a.devices, "wall_s": best, "frames_per_s": a.frames / best,
                          "mpix_per_s": a.frames * w * h / best / 1e6, "host_threads": ncpu,
                          "png_in_mb": in_bytes / 1e6, "tool_says": inner[-1] if inner else "",
                          "pipeline_s": pipe, "pipeline_mpix_per_s": a.frames * w * h / pipe / 1e6 if pipe > 0 else None,
                          "cuda_startup_s": marks.get("dc_create end", 0.0)}), flush=True)
        if a.no_cpu or not ref_available():
            continue
        # ---- CPU arm: the reference's per-index work, frame parallel ----
        ref = RefLib()
        n = a.cpu_frames or a.frames

        def one(i):
            img = cv2.imread(os.path.join(work, f"data{i % a.frames}-depth.png"), cv2.IMREAD_UNCHANGED)
            ok = True
            for j, f in enumerate(fids):
                _, q = ref.cli_frame(f, 0, img, k)
                ok &= cv2.imwrite(os.path.join(work, "out", f"cpu-{j}-{i}.png"), q)
            return ok

        t0 = time.perf_counter()
        with ThreadPoolExecutor(ncpu) as ex:
            assert all(ex.map(one, range(n)))
        dt = time.perf_counter() - t0
        print(json.dumps({"arm": "reference headers + cv2 PNG, frame-parallel", "feature": feat,
                          "size": a.size, "frames": n, "wall_s": dt, "frames_per_s": n / dt,
                          "mpix_per_s": n * w * h / dt / 1e6, "host_threads": ncpu,
                          "cpu": cpu_model}), flush=True)
    if not a.workdir:
        shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()
