#!/bin/bash
# quick GPU check of a kernel change: parity tests, then the bench workloads named in $WLS
set -u
TAG=${1:-q}
WLS=${WLS:-"bearing4_960x540 curvature_1920x1080"}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q ${PYTEST_ARGS:-} 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.txt
cat gpurun_out/${TAG}_pytest.txt
for wl in $WLS; do
  python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-configs --sustain 0 > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_$wl.json"))
    print("$wl", "ms", round(d["ms_per_step"], 4), "MPix/s", round(d["value"]), "frac", round(d["roofline"]["frac"], 3), "e2e", round(d["e2e"]["value"]), d["clocks"])
except Exception as e:
    print("$wl FAILED", e); print(open("gpurun_out/${TAG}_bench_$wl.err").read()[-2000:])
PY
done
