#!/usr/bin/env python
"""Gaussian blur and bilateral pre-filters (row f4): device-resident throughput of
dc_gaussian_blur_device / dc_bilateral_device (CUDA events, f32 frames: 8 B/px) next to
cv2.GaussianBlur / cv2.bilateralFilter on the host cores, frame parallel."""
import ctypes
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import cv2
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_pkg  # noqa: E402

pkg = load_pkg()
w, h, F = 960, 540, 128
dev = torch.device("cuda", 0)
ctx = pkg.Context(0)
ctx.set_frame_size(w, h)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sptr = ctypes.c_void_p(stream.cuda_stream)
d16 = torch.empty((F, h, w), dtype=torch.uint16, device=dev)
ctx.synth_depth_device(0xD2C000B2, 0, F, d16.data_ptr(), sptr)
torch.cuda.synchronize()
d_in = d16.to(torch.float32)
d_out = torch.empty_like(d_in)
host = d_in[:16].cpu().numpy()
ncpu = os.cpu_count()
K = pkg.DC_IN_F32_RANGE
cases = [
    ("gaussian 5x5 sigma 1.5", lambda: ctx._L.dc_gaussian_blur_device(ctx._h, K, d_in.data_ptr(), F, 5, 5, 1.5, 0.0, d_out.data_ptr(), sptr),
     lambda x: cv2.GaussianBlur(x, (5, 5), 1.5), 1e-6),
    ("bilateral d=5 sigma_color 30", lambda: ctx._L.dc_bilateral_device(ctx._h, K, d_in.data_ptr(), F, 5, 30.0, 1.0, d_out.data_ptr(), sptr),
     lambda x: cv2.bilateralFilter(x, 5, 30.0, 1.0), None),
    ("bilateral d=9 sigma_color 30", lambda: ctx._L.dc_bilateral_device(ctx._h, K, d_in.data_ptr(), F, 9, 30.0, 2.0, d_out.data_ptr(), sptr),
     lambda x: cv2.bilateralFilter(x, 9, 30.0, 2.0), None),
]
for name, go, ref, rel in cases:
    for _ in range(3):
        ctx._check(go())
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
    for a, b in evs:
        a.record(stream)
        ctx._check(go())
        b.record(stream)
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in evs) / len(evs)
    want = ref(host[0])
    got = d_out[0].cpu().numpy()
    err = float(np.abs(got - want).max())
    t0 = time.perf_counter()
    with ThreadPoolExecutor(ncpu) as ex:
        list(ex.map(lambda i: ref(host[i % 16]), range(128)))
    cpu_s = time.perf_counter() - t0
    gbs = F * w * h * 8 / (ms * 1e-3) / 1e9
    print(json.dumps({"op": name + " f32", "frames": F, "size": f"{w}x{h}", "ms": ms,
                      "mpix_per_s": F * w * h / (ms * 1e-3) / 1e6, "algorithmic_gb_s": gbs,
                      "max_abs_diff_vs_cv2": err, "value_span": float(host[0].max() - host[0].min()),
                      "cv2_cpu_mpix_per_s": 128 * w * h / cpu_s / 1e6, "cpu_threads": ncpu}))
