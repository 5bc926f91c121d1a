#!/bin/bash
# Rows-per-strip sweep of the row-marching kernels (wave quantisation against halo /
# prologue overhead): DC_ROWS_PER_STRIP (bearing, curvature), DC_FLEX_ROWS, DC_MCE_ROWS.
set -u
mkdir -p gpurun_out
OUT=gpurun_out/${1:-r2_strip_sweep}.txt
: > $OUT
run() {  # env-var value workload
  env $1=$2 python bench.py --workload $3 --steps 30 --warmup 3 --no-cpu-baseline --no-configs --sustain 0 --e2e-frames 8 --e2e-steps 1 > /tmp/sw.json 2> /tmp/sw.err
  python - "$1" "$2" "$3" <<'PY' | tee -a $OUT
import json, sys
try:
    d = json.load(open("/tmp/sw.json"))
    print(sys.argv[3], sys.argv[1], sys.argv[2], "ms", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 4), d["clocks"].get("reasons"))
except Exception as e:
    print(sys.argv[3], sys.argv[1], sys.argv[2], "FAILED", e, open("/tmp/sw.err").read()[-400:])
PY
}
for r in ${ROWS_PAIR:-15 18 20 27 30 36}; do run DC_ROWS_PER_STRIP $r bearing4_960x540; done
for r in ${ROWS_PAIR:-15 18 20 27 30 36}; do run DC_ROWS_PER_STRIP $r curvature_1920x1080; done
for r in ${ROWS_FLEX:-36 45 54 60 72 90}; do run DC_FLEX_ROWS $r flexion_960x540; done
for r in ${ROWS_MCE:-30 36 45 54}; do run DC_MCE_ROWS $r maxcurve_960x540; done
for r in ${ROWS_MCE4K:-30 36 40 45}; do run DC_MCE_ROWS $r range_maxcurve_4096x2160; done
