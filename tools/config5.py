#!/usr/bin/env python
"""BASELINE config 5 on hardware: every depth2x sub-command, PNG in -> PNG out, over
F synthetic 1920x1080 frames in /dev/shm, with --devices D for each D given
(VERDICT r1 item 3). One run per (D, sub-command) under DEPTH2X_PROFILE=1: wall
clock of the process, the pipeline without CUDA start-up (first "dc_create end" ->
last "pipeline drained") and the time spent per stage. On the first D three
output frames per output image kind are compared with the oracle.

    config5.py --frames 1000 --devices 8,4,2,1 --out gpurun_out/r2c5_depth2x.jsonl

The F input files are 8 distinct frames (oracle synthetic depth, the generator of
bench.py) hard-linked F / 8 times: the decoder reads and inflates every one."""
import argparse
import collections
import json
import os
import shutil
import subprocess
import sys
import time

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import (F_BEARING_H, F_FLEXION, F_GAUSS, F_MAX_CURVE, F_MEAN, F_RANGE, Oracle,  # noqa: E402
                        RefLib, kinect_scaled, ref_available)

DEPTH2X = os.path.join(ROOT, "depth-conversions_b200", "bin", "depth2x")
# sub-command -> (arguments, [(output pattern, oracle feature id)])
SUB = collections.OrderedDict([
    ("flexion", [("-o", "out/flexion-{}.png", F_FLEXION)]),
    ("bearing", [("--horizontal", "out/bh-{}.png", F_BEARING_H), ("--vertical", "out/bv-{}.png", F_BEARING_H + 1),
                 ("--diagonal", "out/bd-{}.png", F_BEARING_H + 2),
                 ("--anti-diagonal", "out/ba-{}.png", F_BEARING_H + 3)]),
    ("range", [("-o", "out/range-{}.png", F_RANGE)]),
    ("max-curve", [("-o", "out/mc-{}.png", F_MAX_CURVE)]),
    ("gauss-curvature", [("-o", "out/g-{}.png", F_GAUSS)]),
    ("mean-curvature", [("-o", "out/m-{}.png", F_MEAN)]),
])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--size", default="1920x1080")
    ap.add_argument("--devices", default="8,4,2,1")
    ap.add_argument("--features", default=",".join(SUB))
    ap.add_argument("--workdir", default="/dev/shm/config5")
    ap.add_argument("--out", default="gpurun_out/config5_depth2x.jsonl")
    a = ap.parse_args()
    w, h = (int(x) for x in a.size.split("x"))
    k = kinect_scaled(w, h)
    orc = Oracle()
    chk = RefLib() if ref_available() else orc
    work = a.workdir
    shutil.rmtree(work, ignore_errors=True)
    os.makedirs(os.path.join(work, "out"))
    with open(os.path.join(work, "intrinsic.txt"), "w") as f:
        f.write(f"{w} {h}\n{k['fx']!r} 0.0 {k['cx']!r}\n0.0 {k['fy']!r} {k['cy']!r}\n0.0 0.0 1.0\n")
    t0 = time.perf_counter()
    distinct = [orc.synth_depth(0xD2C00005, i, w, h) for i in range(8)]
    for i in range(a.frames):
        dst = os.path.join(work, f"data{i}-depth.png")
        if i < 8:
            assert cv2.imwrite(dst, distinct[i])
        else:
            os.link(os.path.join(work, f"data{i % 8}-depth.png"), dst)
    in_bytes = sum(os.path.getsize(os.path.join(work, f"data{i}-depth.png")) for i in range(a.frames))
    st = os.statvfs(work)
    out = open(a.out, "a")

    def emit(rec):
        line = json.dumps(rec)
        print(line[:600], flush=True)
        out.write(line + "\n")
        out.flush()

    emit({"what": "inputs", "frames": a.frames, "size": a.size, "png_in_mb": in_bytes / 1e6,
          "generate_s": time.perf_counter() - t0, "shm_free_gb": st.f_bavail * st.f_frsize / 1e9,
          "host_threads": os.cpu_count(), "checker": type(chk).__name__})
    first = True
    for d in (int(x) for x in a.devices.split(",")):
        for feat in a.features.split(","):
            outs = SUB[feat]
            shutil.rmtree(os.path.join(work, "out"), ignore_errors=True)
            os.makedirs(os.path.join(work, "out"))
            cmd = [DEPTH2X, "-c", "intrinsic.txt", "-i", "data{}-depth.png", "-s", "0", "-e", str(a.frames - 1),
                   "--devices", str(d), feat]
            for flag, pat, _ in outs:
                cmd += [flag, pat]
            t0 = time.perf_counter()
            r = subprocess.run(cmd, cwd=work, capture_output=True, text=True, env=dict(os.environ, DEPTH2X_PROFILE="1"))
            wall = time.perf_counter() - t0
            if r.returncode != 0:
                emit({"what": "FAILED", "devices": d, "feature": feat, "rc": r.returncode,
                      "tail": (r.stdout + r.stderr)[-800:]})
                continue
            marks = collections.defaultdict(list)
            for l in r.stderr.splitlines():
                if l.startswith("[profile"):
                    t = float(l[len("[profile"):l.index("s]")])
                    marks[l[l.index("]") + 2:].rsplit(" ", 1)[0]].append(t)
            created = marks.get("dc_create end", [0.0])
            drained = marks.get("pipeline drained", [wall])
            pipe = max(drained) - min(created)
            px = a.frames * w * h
            n_written = len(os.listdir(os.path.join(work, "out")))
            out_bytes = sum(os.path.getsize(os.path.join(work, "out", f)) for f in os.listdir(os.path.join(work, "out")))
            rec = {"what": "run", "devices": d, "feature": feat, "frames": a.frames, "outputs": len(outs),
                   "files_written": n_written, "png_out_mb": out_bytes / 1e6, "wall_s": wall,
                   "wall_mpix_per_s": px / wall / 1e6, "pipeline_s": pipe, "pipeline_mpix_per_s": px / pipe / 1e6,
                   "dc_create_end_s": [round(x, 3) for x in sorted(created)],
                   "first_decoded_s": min(marks.get("decoded", [0.0])),
                   "last_converted_s": max(marks.get("converted", [0.0])),
                   "last_written_s": max(marks.get("encoded+written", [0.0])),
                   "drained_s": [round(x, 3) for x in sorted(drained)],
                   "tool_says": next((l.strip() for l in (r.stdout + r.stderr).splitlines() if "took" in l), "")}
            assert n_written == a.frames * len(outs), (n_written, a.frames, len(outs))
            if first:
                worst = {}
                for _, pat, fid in outs:
                    exact = fid == F_RANGE
                    dmax = 0
                    for i in (0, 3, a.frames - 1):
                        got = cv2.imread(os.path.join(work, pat.format(i)), cv2.IMREAD_UNCHANGED)
                        want = chk.cli_frame(fid, 0, distinct[i % 8], k)[-1]
                        dd = int(np.abs(got.astype(np.int32) - want.astype(np.int32)).max())
                        assert dd <= (0 if exact else 1), (feat, pat, i, dd)
                        dmax = max(dmax, dd)
                    worst[pat] = dmax
                rec["max_lsb_diff_vs_oracle_3_frames"] = worst
            emit(rec)
        first = False
    shutil.rmtree(work, ignore_errors=True)


if __name__ == "__main__":
    main()
