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
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the workload's
    kernel, from the committed ncu --set full capture (profiles/r1_traffic.json)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))[workload]["dram_bytes"]
    except Exception:
        return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                     "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                parts = [x.strip() for x in out.stdout.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            self._halt.wait(0.05)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


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
    sample = f"{n_frames} frames/step of {wl['desc']} (8 distinct synthetic frames, cyclic), in memory, no PNG I/O"
    line = {
        "impl": "reference", "metric": "MPix/s per conversion", "value": value, "unit": "MPix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "desc": wl["desc"], "frames_per_step": n_frames,
                   "width": wl["w"], "height": wl["h"]},
        "cpu_baseline": {"value": value, "unit": "MPix/s", "cores": workers, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "MPix/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


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

    w, h, F = wl["w"], wl["h"], args.frames or wl["frames"]
    px = w * h
    k = kinect_scaled(w, h)
    feature = {"bearing": pkg.DC_BEARING, "flexion": pkg.DC_FLEXION,
               "curvature": pkg.DC_CURVATURE, "max_curve": pkg.DC_MAX_CURVE,
               "range": pkg.DC_RANGE}[wl["feature"]]
    out_kind = pkg.DC_OUT_U16 if wl["out"] == "u16" else pkg.DC_OUT_F32
    out_dtype = torch.uint16 if wl["out"] == "u16" else torch.float32
    out_b = 2 if wl["out"] == "u16" else 4
    n_out = wl["outs"]
    fmap = {"bearing": pkg.DC_BEARING, "flexion": pkg.DC_FLEXION, "curvature": pkg.DC_CURVATURE,
            "max_curve": pkg.DC_MAX_CURVE, "range": pkg.DC_RANGE}
    extra = wl.get("more", [])  # further (feature, outputs) passes over the same frames
    passes = 1 + len(extra)

    ctx = pkg.Context(local)
    ctx.set_pinhole(w, h, k["fx"], k["fy"], k["cx"], k["cy"])
    # A real (non-NULL) stream: NULL would select the context's own stream and
    # the CUDA events below would not bracket the kernels.
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    sptr = ctypes.c_void_p(stream.cuda_stream)

    # frames sharded by frame index, no collective on the data path: rank r owns
    # frames [r*F, (r+1)*F) of every step (depth-conversions_b200/sharding.py)
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "dc_sharding", os.path.join(ROOT, "depth-conversions_b200", "sharding.py"))
    sharding = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sharding)
    my_frames = sharding.rank_frames(rank, world, F)
    d_in = torch.empty((F, h, w), dtype=torch.uint16, device=dev)
    d_out = [torch.empty((F, h, w), dtype=out_dtype, device=dev) for _ in range(n_out)]
    ctx.synth_depth_device(SEED + wl["config"], my_frames.start, len(my_frames), d_in.data_ptr(), sptr)
    job = ctx.make_job(feature, pkg.DC_IN_U16_DEPTH, out_kind, F, d_in.data_ptr(),
                       [o.data_ptr() for o in d_out], clamp=(-20.0, 20.0))
    more_out = [[torch.empty((F, h, w), dtype=out_dtype, device=dev) for _ in range(n)] for _, n in extra]
    more_jobs = [ctx.make_job(fmap[f], pkg.DC_IN_U16_DEPTH, out_kind, F, d_in.data_ptr(),
                              [o.data_ptr() for o in outs], clamp=(-20.0, 20.0))
                 for (f, _), outs in zip(extra, more_out)]
    alg_bytes = F * px * (2 + n_out * out_b) + sum(F * px * (2 + n * out_b) for _, n in extra)
    l2_note = (f"working set per step {alg_bytes / 1e6:.0f} MB in+out "
               f"{'>' if alg_bytes > 126e6 else '<='} 126 MB L2")
    flush = None
    if alg_bytes <= 2 * 126e6:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    kernels = []  # kernel family of every pass, as the library reports it (dc_last_kernel)

    def launch_all():
        ctx.convert_device(job, sptr)
        if len(kernels) < passes:
            kernels.append(ctx.last_kernel)
        for j in more_jobs:
            ctx.convert_device(j, sptr)
            if len(kernels) < passes:
                kernels.append(ctx.last_kernel)

    def step():
        if flush is not None:
            flush.fill_(1)
        launch_all()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    # --- timed region: K steps, CUDA events on the launching stream ----------
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    launches1 = ctx.kernel_launches
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in evs:
        if flush is not None:
            flush.fill_(1)
        a.record(stream)
        launch_all()
        b.record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches_timed = ctx.kernel_launches - launches1
    kern_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = float(sum(kern_ms))
    if flush is None:
        # no flush kernels between steps: the wall bracket is the honest step time
        total_ms = max(total_ms, 0.0)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    # MPix/s per conversion: input pixels of every pass (a pass = one depth2x sub-command)
    value = world * args.steps * F * px * passes / (total_ms * 1e-3) / 1e6

    # --- e2e: host buffers through dc_convert_host ---------------------------
    Fe = min(args.e2e_frames or 128, F)
    h_in = torch.empty((Fe, h, w), dtype=torch.uint16).pin_memory()
    h_in.copy_(d_in[:Fe].cpu())
    h_out = [torch.empty((Fe, h, w), dtype=out_dtype).pin_memory() for _ in range(n_out)]
    ejob = ctx.make_job(feature, pkg.DC_IN_U16_DEPTH, out_kind, Fe, h_in.data_ptr(),
                        [o.data_ptr() for o in h_out], clamp=(-20.0, 20.0))
    more_hout = [[torch.empty((Fe, h, w), dtype=out_dtype).pin_memory() for _ in range(n)] for _, n in extra]
    ejobs = [ejob] + [ctx.make_job(fmap[f], pkg.DC_IN_U16_DEPTH, out_kind, Fe, h_in.data_ptr(),
                                   [o.data_ptr() for o in outs], clamp=(-20.0, 20.0))
                      for (f, _), outs in zip(extra, more_hout)]
    e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        for j in ejobs:
            ctx._check(ctx.convert_host_raw(j))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e_steps):
        for j in ejobs:
            ctx._check(ctx.convert_host_raw(j))
    torch.cuda.synchronize()
    e_sec = time.perf_counter() - t0
    te = torch.tensor([e_sec], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e_value = world * e_steps * Fe * px * passes / float(te.item()) / 1e6
    clocks = sampler.stop() if sampler else None  # sampled from the warm-up to the end of e2e
    # parity spot check of the e2e result against the device-resident result
    same = all(bool(torch.equal(h_out[i][:1], d_out[i][:1].cpu())) for i in range(n_out))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    avg_kernel_ms = float(np.mean(kern_ms))
    achieved = alg_bytes / (avg_kernel_ms * 1e-3) / 1e9
    line = {
        "metric": "MPix/s per conversion", "value": value, "unit": "MPix/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "desc": wl["desc"], "frames_per_step_per_gpu": F,
                   "width": w, "height": h, "outputs": n_out, "out_type": wl["out"],
                   "l2": l2_note + ("; 256 MB L2 flush write between steps" if flush is not None else
                                    "; inputs+outputs larger than L2, no flush needed"),
                   "sharding": "frame index, no collective"},
        "e2e": {"value": e_value, "unit": "MPix/s", "h2d_bytes_per_step": Fe * px * 2 * passes,
                "d2h_bytes_per_step": Fe * px * (n_out + sum(n for _, n in extra)) * out_b,
                "frames_per_step_per_gpu": Fe,
                "steps": e_steps, "api": "dc_convert_host (pinned host buffers)",
                "matches_device_result": same},
        "gpu_launches": int(launches_timed),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(args.workload), "peak_source": peak_src,
                     "kernel": " + ".join(f"{k} [{f}, u16 depth -> {wl['out']}]" for k, f in
                                          zip(kernels, [wl["feature"]] + [f for f, _ in extra])),
                     "launch_ms": avg_kernel_ms,
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "bytes_per_pixel": alg_bytes / (F * px * passes), "passes_per_step": passes},
        "clocks": clocks,
        "wall_ms_per_step": 1e3 * t_wall / args.steps,
    }
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        n_cpu = cpu_sample_frames(wl, 12.0, cores)
        dt, workers, kind = cpu_reference_run(wl, n_cpu)
        line["cpu_baseline"] = {
            "value": n_cpu * px * passes / dt / 1e6, "unit": "MPix/s", "cores": workers, "kind": kind,
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
    ap.add_argument("--e2e-frames", type=int, default=0)
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
