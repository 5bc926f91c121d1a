#!/bin/bash
# e2e (dc_convert_host, pinned host buffers) against the staging chunk size
set -u
mkdir -p gpurun_out
for mb in ${CHUNKS:-4 8 16 24 48 96}; do
  for wl in ${WLS:-flexion_960x540 bearing4_960x540 curvature_1920x1080 range_maxcurve_4096x2160}; do
    DC_CHUNK_MB=$mb python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/chunk_${mb}_$wl.json 2>/dev/null
    python - <<PY
import json
d = json.load(open("gpurun_out/chunk_${mb}_$wl.json"))
e = d["e2e"]
print("chunk ${mb} MB $wl e2e", round(e["value"]), "MPix/s", "H2D+D2H GB/s", round((e["h2d_bytes_per_step"] + e["d2h_bytes_per_step"]) * e["value"] * 1e6 / (e["frames_per_step_per_gpu"] * d["config"]["width"] * d["config"]["height"]) / 1e9 / max(1, d["roofline"]["passes_per_step"]), 1))
PY
  done
done | tee gpurun_out/chunk_summary.txt
