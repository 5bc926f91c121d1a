#!/usr/bin/env python
"""Median pre-filter (row f4): device-resident throughput of dc_median_device
(CUDA events) next to cv2.medianBlur on the host cores, frame parallel."""
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
w, h, F = 960, 540, 256
dev = torch.device("cuda", 0)
ctx = pkg.Context(0)
ctx.set_frame_size(w, h)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
sptr = ctypes.c_void_p(stream.cuda_stream)
d_in = torch.empty((F, h, w), dtype=torch.uint16, device=dev)
d_out = torch.empty_like(d_in)
ctx.synth_depth_device(0xD2C000B0, 0, F, d_in.data_ptr(), sptr)
host = d_in[:32].cpu().numpy()
ncpu = os.cpu_count()
for k in (3, 5):
    def go():
        ctx._check(ctx._L.dc_median_device(ctx._h, pkg.DC_IN_U16_RANGE, d_in.data_ptr(), F, k,
                                           d_out.data_ptr(), sptr))
    for _ in range(3):
        go()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for a, b in evs:
        a.record(stream)
        go()
        b.record(stream)
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in evs) / len(evs)
    assert np.array_equal(d_out[0].cpu().numpy(), cv2.medianBlur(host[0], k))
    t0 = time.perf_counter()
    with ThreadPoolExecutor(ncpu) as ex:
        list(ex.map(lambda i: cv2.medianBlur(host[i % 32], k), range(256)))
    cpu_s = time.perf_counter() - t0
    gbs = F * w * h * 4 / (ms * 1e-3) / 1e9
    print(json.dumps({"op": f"median {k}x{k} u16", "frames": F, "size": f"{w}x{h}", "ms": ms,
                      "mpix_per_s": F * w * h / (ms * 1e-3) / 1e6, "algorithmic_gb_s": gbs,
                      "frac_of_6537": gbs / 6537.0,
                      "cv2_cpu_mpix_per_s": 256 * w * h / cpu_s / 1e6, "cpu_threads": ncpu}))
