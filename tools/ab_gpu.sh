#!/bin/bash
# A/B timing of kernel variants: for each lib in $LIBS run the workloads in $WLS (no parity check)
set -u
TAG=${1:-ab}
WLS=${WLS:-"bearing4_960x540 curvature_1920x1080"}
mkdir -p gpurun_out
for lib in $LIBS; do
 for wl in $WLS; do
  DC_LIB=$PWD/depth-conversions_b200/variants/libdc_$lib.so python bench.py --workload $wl --steps 30 --warmup 3 --no-cpu-baseline --no-configs --sustain 0 ${BENCH_ARGS:-} > gpurun_out/${TAG}_${lib}_$wl.json 2> gpurun_out/${TAG}_${lib}_$wl.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_${lib}_$wl.json"))
    print("$lib $wl", "ms", round(d["ms_per_step"], 4), "frac", round(d["roofline"]["frac"], 3), "match", d["e2e"].get("matches_device_result"), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print("$lib $wl FAILED", e); print(open("gpurun_out/${TAG}_${lib}_$wl.err").read()[-1500:])
PY
 done
done | tee gpurun_out/${TAG}_summary.txt
