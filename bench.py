#!/usr/bin/env python
"""bench.py -- throughput of the depth -> feature hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload NAME] [--frames F]

A "step" is one pass of the hot path over one batch of F synthetic frames per
GPU: one fused kernel launch through the C ABI (dc_convert_device). Prints ONE
JSON line (rank 0). Default workload = BASELINE.json configs[1]:
`depth2x bearing` (horizontal/vertical/diagonal/antidiagonal) on 960x540
pinhole depth, u16 in -> 4 x u16 out.

    value     MPix/s (input pixels converted per second, whole job), inputs
              resident in HBM, CUDA-event timed, max over ranks
    e2e       the same through dc_convert_host with pinned HOST buffers: H2D of
              the frames and D2H of every output inside the timed region
    roofline  algorithmic bytes (each input pixel read once + each output pixel
              written once; per-camera tables count 0) / kernel time, against
              the measured HBM copy peak in MEASURED_PEAKS.json
    cpu_baseline  the reference's taskflow frame-parallel CPU path
              (oracle/_ref: the reference's own headers compiled; else the C
              restatement) on a bounded sample, on the host cores of this box

    sustained the timed region's launches back to back for >= 2.5 s, with the SM
              clock, power and throttle reasons sampled INSIDE that window
    configs   (N = 1) the other BASELINE.json configs and the library API's f32 forms,
              device resident: MPix/s, ms per step, kernel, roofline fraction each

--impl reference times that CPU path alone (rank 0 only under torchrun).
PyTorch is plumbing here: device buffers, the stream, CUDA events, the
torch.distributed barrier / max-reduce. The kernels are this repo's.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SEED = 0xD2C00000

# name -> (w, h, feature, in_kind, out_kind, n_outputs, default frames/step,
#          oracle feature ids, description)
WORKLOADS = {
    # BASELINE.json configs[1] (the headline)
    "bearing4_960x540": dict(w=960, h=540, feature="bearing", out="u16", outs=4, frames=256,
                             ref_features=[1, 2, 3, 4], config=2,
                             desc="depth2x bearing h/v/d/a, 960x540 pinhole depth, u16 -> 4x u16"),
    # configs[0]
    "flexion_960x540": dict(w=960, h=540, feature="flexion", out="u16", outs=1, frames=256,
                            ref_features=[5], config=1,
                            desc="depth2x flexion --type pinhole-depth, 960x540, u16 -> u16"),
    # configs[2]
    "curvature_1920x1080": dict(w=1920, h=1080, feature="curvature", out="f32", outs=2, frames=64,
                                ref_features=[6, 7], config=3,
                                desc="Gaussian + mean curvature, 1920x1080 pinhole depth, u16 -> 2x f32"),
    # configs[3]
    "range_maxcurve_4096x2160": dict(w=4096, h=2160, feature="max_curve", out="u16", outs=2,
                                     frames=16, ref_features=[0, 8], config=4,
                                     desc="laserscan range fused with max-curve, 4096x2160, u16 -> 2x u16"),
    # the library API's types: f32 range image in, f32 out
    "bearing4_f32_960x540": dict(w=960, h=540, feature="bearing", out="f32", outs=4, frames=256,
                                 inp="f32", ref_features=[1, 2, 3, 4], config=2,
                                 desc="depth_to_bearing x4 over math::image<float>, 960x540, f32 range -> 4x f32"),
    "flexion_f32_960x540": dict(w=960, h=540, feature="flexion", out="f32", outs=1, frames=256,
                                inp="f32", ref_features=[5], config=1,
                                desc="depth_to_flexion over math::image<float>, 960x540, f32 range -> f32"),
    # Real = double (what src/unittest/conversion/*.cpp instantiate)
    "bearing4_f64_960x540": dict(w=960, h=540, feature="bearing", out="f64", outs=4, frames=64,
                                 ref_features=[1, 2, 3, 4], config=2,
                                 desc="depth_to_bearing<double> x4, 960x540 pinhole depth, u16 -> 4x f64"),
    "flexion_f64_960x540": dict(w=960, h=540, feature="flexion", out="f64", outs=1, frames=64,
                                ref_features=[5], config=1,
                                desc="depth_to_flexion<double>, 960x540 pinhole depth, u16 -> f64"),
    "range_960x540": dict(w=960, h=540, feature="range", out="u16", outs=1, frames=256,
                          ref_features=[0], config=0,
                          desc="depth2x range, 960x540 pinhole depth, u16 -> u16"),
    "maxcurve_960x540": dict(w=960, h=540, feature="max_curve", out="u16", outs=1, frames=256,
                             ref_features=[8], config=0,
                             desc="depth2x max-curve, 960x540 pinhole depth, u16 -> u16"),
    # configs[4], kernel-only form: every feature type over the same 1080p frames (four jobs
    # per step: bearing x4, flexion, Gaussian + mean curvature, range + max-curve; all u16)
    "all_1920x1080": dict(w=1920, h=1080, feature="bearing", out="u16", outs=4, frames=64,
                          more=[("flexion", 1), ("curvature", 2), ("max_curve", 2)],
                          ref_features=[1, 2, 3, 4, 5, 6, 7, 0, 8], config=5,
                          desc="all depth2x feature types, 1920x1080 pinhole depth, u16 -> 9x u16 "
                               "(bearing x4 | flexion | gauss + mean | range + max-curve: 4 passes)"),
}


def kinect_scaled(w, h):
    f = np.float32
    return dict(w=w, h=h, fx=float(f(519.226 * w / 960)), fy=float(f(522.23 * h / 540)),
                cx=float(f(479.462 * w / 960)), cy=float(f(272.737 * h / 540)))


def ncu_traffic(workload):
    """(dram__bytes_read.sum + dram__bytes_write.sum of one launch of the workload's
    kernel, where that number comes from): NOT measured in this run -- it needs ncu --
    but read from the committed summary of the ncu --set full capture of the same
    command (profiles/r2_traffic.json, else the round-1 file)."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            e = json.load(open(os.path.join(ROOT, "profiles", name)))[workload]
            return e["dram_bytes"], f"profiles/{name}: {e.get('source', '')} [{e.get('kernel', '')}]"
        except Exception:
            continue
    return None, None


def box_dma_ceiling(n_gpus, h2d, d2h):
    """Host<->device copy ceiling of the box with n_gpus GPUs copying at once, from the
    committed probe (profiles/r2_pcie_probe.json; NOT measured in this run): the
    bearing-like 1:4 mix when both directions carry data, else the one-direction row."""
    try:
        e = json.load(open(os.path.join(ROOT, "profiles", "r2_pcie_probe.json")))
        row = e["gpus"][str(n_gpus)]
        key = "h2d+d2h(1:4)" if h2d and d2h else ("h2d" if h2d else "d2h")
        return row[key], f"profiles/r2_pcie_probe.json [{key}, {n_gpus} GPU(s) active]"
    except Exception:
        return None, None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons / power, time-stamped, every 25 ms; the
    summary of a window is taken from the samples that fall inside it."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self._halt = index, [], threading.Event()

    def run(self):
        # one long-lived nvidia-smi in loop mode: a sample every 25 ms without a
        # process start per sample
        try:
            proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "25"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            return
        self._proc = proc
        for line in proc.stdout:
            parts = [x.strip() for x in line.strip().split(",")]
            if len(parts) >= 7:
                self.samples.append((time.perf_counter(), parts))
            if self._halt.is_set():
                break

    def stop(self):
        self._halt.set()
        try:
            self._proc.terminate()
        except Exception:
            pass
        self.join(timeout=6)

    def summary(self, t0=None, t1=None):
        sel = [p for t, p in self.samples if (t0 is None or t >= t0) and (t1 is None or t <= t1)]
        num = lambda v: float(v) if v.replace(".", "").isdigit() else None
        sm = [num(p[0]) for p in sel if num(p[0]) is not None]
        mx = [num(p[1]) for p in sel if num(p[1]) is not None]
        pw = [num(p[6]) for p in sel if num(p[6]) is not None]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for p in sel for n, v in zip(names, p[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_mhz_min": min(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "power_w": float(np.median(pw)) if pw else None, "samples": len(sel)}


def cpu_reference_run(wl, n_frames, threads=0):
    """The reference's frame-parallel CPU path on `n_frames` frames (cyclic over
    8 distinct synthetic frames). Returns (seconds, workers, kind)."""
    from oracle_lib import Oracle, RefLib, ref_available
    orc = Oracle()
    w, h = wl["w"], wl["h"]
    k = kinect_scaled(w, h)
    distinct = np.stack([orc.synth_depth(SEED + wl["config"], f, w, h) for f in range(8)])
    if ref_available():
        ref = RefLib()
        ref.cli_batch(wl["ref_features"], 0, distinct[:1], k, want_outputs=False, n_frames=1)
        t0 = time.perf_counter()
        _, workers = ref.cli_batch(wl["ref_features"], 0, distinct, k, threads=threads,
                                   want_outputs=False, n_frames=n_frames)
        return time.perf_counter() - t0, workers, "reference"
    # oracle port: scalar, one thread
    t0 = time.perf_counter()
    for i in range(n_frames):
        for f in wl["ref_features"]:
            orc.cli_frame(f, 0, distinct[i % 8], k)
    return time.perf_counter() - t0, 1, "port"


def cpu_sample_frames(wl, seconds_target, cores):
    """Frames for a bounded CPU sample of about `seconds_target` seconds."""
    # measured here: ~0.07 s per 960x540 frame per bearing direction per core;
    # max-curve ~0.4 s, flexion 0.12 s, curvature ~0.08 s, range ~0.01 s
    per_px = {0: 0.02e-6, 1: 0.13e-6, 2: 0.13e-6, 3: 0.13e-6, 4: 0.13e-6, 5: 0.25e-6,
              6: 0.16e-6, 7: 0.18e-6, 8: 0.8e-6}
    sec_per_frame = sum(per_px[f] for f in wl["ref_features"]) * wl["w"] * wl["h"] + 0.02e-6 * wl["w"] * wl["h"]
    n = int(seconds_target * max(1, cores) / sec_per_frame)
    return max(cores, min(n, 200000))


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_frames = cpu_sample_frames(wl, 6.0, cores)
    px = wl["w"] * wl["h"]
    for _ in range(args.warmup):
        cpu_reference_run(wl, max(cores, n_frames // 8))
    t_total, workers, kind = 0.0, cores, "reference"
    for _ in range(args.steps):
        dt, workers, kind = cpu_reference_run(wl, n_frames)
        t_total += dt
    passes = 1 + len(wl.get("more", []))
    value = args.steps * n_frames * px * passes / t_total / 1e6
    sample = (f"{n_frames} frames per step (a bounded sample of the workload in `config`: "
              f"8 distinct synthetic frames, cyclic), in memory, no PNG I/O")
    line = {
        "impl": "reference", "metric": "MPix/s per conversion", "value": value, "unit": "MPix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, wl, args.frames or wl["frames"]),
        "cpu_baseline": {"value": value, "unit": "MPix/s", "cores": workers, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "MPix/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


OUT_BYTES = {"u16": 2, "f32": 4, "f64": 8}


def algorithmic_bytes(wl, F):
    """Each input pixel read once + each output pixel written once."""
    px = wl["w"] * wl["h"]
    in_b = 4 if wl.get("inp") == "f32" else 2
    out_b = OUT_BYTES[wl["out"]]
    total = F * px * (in_b + wl["outs"] * out_b)
    for _, n in wl.get("more", []):
        total += F * px * (in_b + n * out_b)
    return total


def config_dict(args, wl, F):
    """The same dictionary for both arms (--impl ours / reference): the workload."""
    alg = algorithmic_bytes(wl, F)
    return {"workload": args.workload, "desc": wl["desc"], "frames_per_step_per_gpu": F,
            "width": wl["w"], "height": wl["h"], "outputs": wl["outs"], "in_type": wl.get("inp", "u16"),
            "out_type": wl["out"],
            "l2": (f"working set per step {alg / 1e6:.0f} MB in+out "
                   f"{'>' if alg > 126e6 else '<='} 126 MB L2"
                   + ("; 256 MB L2 flush write between steps" if alg <= 2 * 126e6 else
                      "; inputs+outputs larger than L2, no flush needed")),
            "sharding": "frame index, no collective"}


class Workload:
    """Device-resident buffers and jobs of one workload on one GPU."""

    def __init__(self, pkg, torch, dev, local, stream, wl, F, rank, world, sharding):
        self.pkg, self.torch, self.wl, self.F, self.stream = pkg, torch, wl, F, stream
        self.sptr = ctypes.c_void_p(stream.cuda_stream)
        w, h = wl["w"], wl["h"]
        self.px = w * h
        k = kinect_scaled(w, h)
        fmap = {"bearing": pkg.DC_BEARING, "flexion": pkg.DC_FLEXION, "curvature": pkg.DC_CURVATURE,
                "max_curve": pkg.DC_MAX_CURVE, "range": pkg.DC_RANGE}
        self.out_kind = {"u16": pkg.DC_OUT_U16, "f32": pkg.DC_OUT_F32, "f64": pkg.DC_OUT_F64}[wl["out"]]
        self.out_dtype = {"u16": torch.uint16, "f32": torch.float32, "f64": torch.float64}[wl["out"]]
        self.extra = wl.get("more", [])
        self.passes = 1 + len(self.extra)
        self.features = [wl["feature"]] + [f for f, _ in self.extra]
        self.ctx = pkg.Context(local)
        self.ctx.set_pinhole(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
        frames = sharding.rank_frames(rank, world, F)
        d_u16 = torch.empty((F, h, w), dtype=torch.uint16, device=dev)
        self.ctx.synth_depth_device(SEED + wl["config"], frames.start, len(frames), d_u16.data_ptr(), self.sptr)
        if wl.get("inp") == "f32":
            # the library API's input: the range image as f32 (depth_to_laserscan of the frames)
            self.d_in = torch.empty((F, h, w), dtype=torch.float32, device=dev)
            self.ctx.convert_device(self.ctx.make_job(pkg.DC_RANGE, pkg.DC_IN_U16_DEPTH, pkg.DC_OUT_F32, F,
                                                      d_u16.data_ptr(), [self.d_in.data_ptr()]), self.sptr)
            torch.cuda.synchronize()
            del d_u16
            self.in_kind = pkg.DC_IN_F32_RANGE
        else:
            self.d_in, self.in_kind = d_u16, pkg.DC_IN_U16_DEPTH
        self.d_out = [[torch.empty((F, h, w), dtype=self.out_dtype, device=dev) for _ in range(n)]
                      for n in [wl["outs"]] + [n for _, n in self.extra]]
        self.jobs = [self.ctx.make_job(fmap[f], self.in_kind, self.out_kind, F, self.d_in.data_ptr(),
                                       [o.data_ptr() for o in outs], clamp=(-20.0, 20.0))
                     for f, outs in zip(self.features, self.d_out)]
        self.alg_bytes = algorithmic_bytes(wl, F)
        self.flush = (torch.empty(256 << 20, dtype=torch.uint8, device=dev)
                      if self.alg_bytes <= 2 * 126e6 else None)
        self.kernels = []

    def launch_all(self):
        for j in self.jobs:
            self.ctx.convert_device(j, self.sptr)
            if len(self.kernels) < self.passes:
                self.kernels.append(self.ctx.last_kernel)

    def timed(self, steps, warmup, barrier):
        """K steps, CUDA events on the launching stream around the kernels of each step
        (the L2 flush, where one runs, sits between the brackets). Returns the per-step
        kernel times in ms and the kernel launches counted inside the region."""
        torch = self.torch
        for _ in range(warmup):
            if self.flush is not None:
                self.flush.fill_(1)
            self.launch_all()
        barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
               for _ in range(steps)]
        launches0 = self.ctx.kernel_launches
        barrier()
        t0 = time.perf_counter()
        for a, b in evs:
            if self.flush is not None:
                self.flush.fill_(1)
            a.record(self.stream)
            self.launch_all()
            b.record(self.stream)
        barrier()
        wall = time.perf_counter() - t0
        return [a.elapsed_time(b) for a, b in evs], self.ctx.kernel_launches - launches0, wall, t0

    def sustained(self, seconds):
        """Back-to-back launches for at least `seconds` (no flush: the batch is larger
        than L2 or the same flush-free loop a long job runs). One event pair around all."""
        torch = self.torch
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n, t0 = 0, time.perf_counter()
        a.record(self.stream)
        while True:
            for _ in range(16):
                self.launch_all()
            n += 16
            if time.perf_counter() - t0 >= seconds:
                break
            if n % 512 == 0:
                self.stream.synchronize()  # bound the launch queue
        b.record(self.stream)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        return n, a.elapsed_time(b), t0, t1

    def roofline(self, avg_ms, peak, peak_src, name):
        achieved = self.alg_bytes / (avg_ms * 1e-3) / 1e9
        traffic, tsrc = ncu_traffic(name)
        return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": tsrc,
                "peak_source": peak_src,
                "kernel": " + ".join(f"{k} [{f}, {self.wl.get('inp', 'u16')} -> {self.wl['out']}]"
                                     for k, f in zip(self.kernels, self.features)),
                "launch_ms": avg_ms, "algorithmic_bytes_per_launch": self.alg_bytes,
                "bytes_per_pixel": self.alg_bytes / (self.F * self.px * self.passes),
                "passes_per_step": self.passes}

    def close(self):
        self.ctx.close()


def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    from __graft_entry__ import load_pkg
    pkg = load_pkg()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local))
    if args.gpus != world:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # A real (non-NULL) stream: NULL would select the context's own stream and
    # the CUDA events below would not bracket the kernels.
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0

    # frames sharded by frame index, no collective on the data path: rank r owns
    # frames [r*F, (r+1)*F) of every step (depth-conversions_b200/sharding.py)
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "dc_sharding", os.path.join(ROOT, "depth-conversions_b200", "sharding.py"))
    sharding = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sharding)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    w, h, F = wl["w"], wl["h"], args.frames or wl["frames"]
    px = w * h
    W = Workload(pkg, torch, dev, local, stream, wl, F, rank, world, sharding)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()

    # --- the timed region -----------------------------------------------------
    kern_ms, launches_timed, t_wall, t_region0 = W.timed(args.steps, args.warmup, barrier)
    t_region1 = time.perf_counter()
    total_ms = max_over_ranks(float(sum(kern_ms)))
    # MPix/s per conversion: input pixels of every pass (a pass = one depth2x sub-command)
    value = world * args.steps * F * px * W.passes / (total_ms * 1e-3) / 1e6

    # --- sustained: the same launches back to back for >= 2 s, clocks sampled inside
    sustained = None
    if args.sustain > 0:
        barrier()
        n_s, ms_s, ts0, ts1 = W.sustained(args.sustain)
        ms_s = max_over_ranks(ms_s)
        sustained = {"seconds": ms_s * 1e-3, "launches": n_s * W.passes,
                     "value": world * n_s * F * px * W.passes / (ms_s * 1e-3) / 1e6, "unit": "MPix/s",
                     "ms_per_step": ms_s / n_s, "window": (ts0, ts1)}

    # --- e2e: host buffers through dc_convert_host, the frames and steps of the main leg
    Fe = min(args.e2e_frames or F, F)
    out_b = OUT_BYTES[wl["out"]]
    in_b = 4 if wl.get("inp") == "f32" else 2
    h_in = torch.empty((Fe, h, w), dtype=W.d_in.dtype).pin_memory()
    h_in.copy_(W.d_in[:Fe].cpu())
    h_outs = [[torch.empty((Fe, h, w), dtype=W.out_dtype).pin_memory() for _ in outs] for outs in W.d_out]
    fmap = {"bearing": pkg.DC_BEARING, "flexion": pkg.DC_FLEXION, "curvature": pkg.DC_CURVATURE,
            "max_curve": pkg.DC_MAX_CURVE, "range": pkg.DC_RANGE}
    ejobs = [W.ctx.make_job(fmap[f], W.in_kind, W.out_kind, Fe, h_in.data_ptr(),
                            [o.data_ptr() for o in outs], clamp=(-20.0, 20.0))
             for f, outs in zip(W.features, h_outs)]
    e_steps = args.e2e_steps or args.steps
    for _ in range(2):
        for j in ejobs:
            W.ctx._check(W.ctx.convert_host_raw(j))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e_steps):
        for j in ejobs:
            W.ctx._check(W.ctx.convert_host_raw(j))
    torch.cuda.synchronize()
    e_sec = max_over_ranks(time.perf_counter() - t0)
    e_value = world * e_steps * Fe * px * W.passes / e_sec / 1e6
    # parity spot check of the e2e result against the device-resident result
    same = all(bool(torch.equal(ho[:1], do[:1].cpu()))
               for houts, douts in zip(h_outs, W.d_out) for ho, do in zip(houts, douts))
    n_out_all = sum(len(o) for o in W.d_out)
    # The same bytes as bare copies, measured here and now on this box with all ranks copying
    # together (no kernel, no staging logic): one cudaMemcpyAsync per buffer (torch copy_ of
    # pinned <-> device tensors), uploads on one stream, downloads on another. This is the
    # end-to-end ceiling the box's DMA fabric sets for this step.
    up, down = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    def bare_copies():
        with torch.cuda.stream(up):
            for _ in range(W.passes):
                W.d_in[:Fe].copy_(h_in, non_blocking=True)
        with torch.cuda.stream(down):
            for houts, douts in zip(h_outs, W.d_out):
                for ho, do in zip(houts, douts):
                    ho.copy_(do[:Fe], non_blocking=True)
    saved_in = W.d_in[:Fe].clone()
    bare_copies()
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    p_steps = max(2, min(e_steps, 10))
    for _ in range(p_steps):
        bare_copies()
    torch.cuda.synchronize()
    p_sec = max_over_ranks(time.perf_counter() - t0)
    W.d_in[:Fe].copy_(saved_in)
    del saved_in
    if sampler:
        sampler.stop()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    avg_kernel_ms = float(np.mean(kern_ms))
    line = {
        "metric": "MPix/s per conversion", "value": value, "unit": "MPix/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64" if wl["out"] == "f64" else "f32", "data": "synthetic",
        "config": config_dict(args, wl, F),
        "e2e": {"value": e_value, "unit": "MPix/s", "h2d_bytes_per_step": Fe * px * in_b * W.passes,
                "d2h_bytes_per_step": Fe * px * n_out_all * out_b,
                "frames_per_step_per_gpu": Fe, "steps": e_steps,
                "api": "dc_convert_host (pinned host buffers)", "matches_device_result": same},
        "gpu_launches": int(launches_timed),
        "roofline": W.roofline(avg_kernel_ms, peak, peak_src, args.workload),
        # clocks: every nvidia-smi sample (25 ms period) from the first timed step to the end
        # of the sustained window; the K-step region alone lasts total_ms (often < 25 ms)
        "clocks": dict(sampler.summary(t_region0, sustained["window"][1] if sustained else t_region1),
                       timed_region_ms=total_ms),
        "wall_ms_per_step": 1e3 * t_wall / args.steps,
    }
    # e2e against what the box's host<->device DMA delivers to n GPUs copying together
    e = line["e2e"]
    step_bytes = e["h2d_bytes_per_step"] + e["d2h_bytes_per_step"]
    moved = world * e_steps * step_bytes / e_sec / 1e9
    bare = world * p_steps * step_bytes / p_sec / 1e9
    e.update(moved_gbs=moved, box_dma_gbs=bare, frac_of_box_dma=moved / bare,
             box_dma_source=f"bare pinned cudaMemcpyAsync of the same buffers, {world} GPU(s) copying at once, "
                            f"{p_steps} steps, measured in this run")
    dma, dma_src = box_dma_ceiling(world, e["h2d_bytes_per_step"], e["d2h_bytes_per_step"])
    if dma:  # the stand-alone probe on the pool's 8-GPU box (boxes of the pool differ)
        e.update(probe_8gpu_box_gbs=dma, probe_source=dma_src)
    if sustained:
        win = sustained.pop("window")
        sustained["clocks"] = sampler.summary(*win)
        sustained["roofline_frac"] = (W.alg_bytes / (sustained["ms_per_step"] * 1e-3) / 1e9) / peak
        sustained["note"] = ("the timed region's launches repeated back to back without the L2 flush; "
                             "clocks = nvidia-smi samples inside this window")
        line["sustained"] = sustained
    W.close()

    # --- the other BASELINE configs, device resident (N = 1 only): one line for the driver
    if world == 1 and not args.no_configs:
        configs = {}
        names = [n for n in ("flexion_960x540", "bearing4_960x540", "curvature_1920x1080",
                             "range_maxcurve_4096x2160", "all_1920x1080", "bearing4_f32_960x540",
                             "flexion_f32_960x540", "bearing4_f64_960x540", "flexion_f64_960x540")
                 if n != args.workload]
        for name in names:
            wl2 = WORKLOADS[name]
            X = Workload(pkg, torch, dev, local, stream, wl2, wl2["frames"], 0, 1, sharding)
            ms, _, _, _ = X.timed(10, 3, barrier)
            avg = float(np.mean(ms))
            rf = X.roofline(avg, peak, peak_src, name)
            configs[name] = {"baseline_config": wl2["config"], "desc": wl2["desc"],
                             "frames_per_step": wl2["frames"],
                             "value": wl2["frames"] * wl2["w"] * wl2["h"] * X.passes / (avg * 1e-3) / 1e6,
                             "unit": "MPix/s", "ms_per_step": avg, "steps": 10, "warmup": 3,
                             "kernel": rf["kernel"], "bytes_per_pixel": rf["bytes_per_pixel"],
                             "algorithmic_bytes_per_launch": rf["algorithmic_bytes_per_launch"],
                             "roofline_frac": rf["frac"], "achieved_gbs": rf["achieved"],
                             "traffic": rf["traffic"], "traffic_source": rf["traffic_source"],
                             "l2_flush": X.flush is not None}
            X.close()
            del X
            torch.cuda.empty_cache()
        line["configs"] = configs
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_cpu = cpu_sample_frames(wl, 12.0, cores)
        dt, workers, kind = cpu_reference_run(wl, n_cpu)
        line["cpu_baseline"] = {
            "value": n_cpu * px * W.passes / dt / 1e6, "unit": "MPix/s", "cores": workers, "kind": kind,
            "sample": f"{n_cpu} frames of the same workload (8 distinct synthetic frames, cyclic), "
                      f"taskflow frame-parallel, in memory, {dt:.1f} s"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="bearing4_960x540", choices=sorted(WORKLOADS))
    ap.add_argument("--frames", type=int, default=0, help="frames per step per GPU")
    ap.add_argument("--e2e-frames", type=int, default=0, help="frames per e2e step (0: as the main leg)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="e2e steps (0: as the main leg)")
    ap.add_argument("--sustain", type=float, default=2.5,
                    help="seconds of back-to-back launches with clocks sampled inside (0: off)")
    ap.add_argument("--no-configs", action="store_true",
                    help="skip the device-resident lines of the other BASELINE configs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
