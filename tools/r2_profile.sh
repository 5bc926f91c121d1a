#!/bin/bash
# Round 2 measurement at HEAD, one GPU: the GPU test suite, the default bench line
# (configs + sustained window + CPU baseline), the reference arm, the ncu launch
# list of the default command and one --set full capture per hot kernel (raw and
# source pages as CSV; the reports themselves are too large to travel back).
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -40 > gpurun_out/${TAG}_pytest_gpu.txt
tail -3 gpurun_out/${TAG}_pytest_gpu.txt
python bench.py > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err
tail -c 3500 gpurun_out/${TAG}_bench_default.json; tail -3 gpurun_out/${TAG}_bench_default.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bearing4_960x540.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --sustain 0 > gpurun_out/${TAG}_ncu_launches.log 2>&1
for wl in bearing4_960x540 flexion_960x540 curvature_1920x1080 range_maxcurve_4096x2160 maxcurve_960x540 bearing4_f32_960x540 flexion_f32_960x540 range_960x540; do
  python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-configs --sustain 0 > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_$wl.json"))
    print("$wl", "ms", round(d["ms_per_step"], 4), "MPix/s", round(d["value"]), "frac", round(d["roofline"]["frac"], 3), d["roofline"]["kernel"], "e2e", round(d["e2e"]["value"]))
except Exception as e:
    print("$wl FAILED", e); print(open("gpurun_out/${TAG}_bench_$wl.err").read()[-1500:])
PY
  timeout 300 ncu --set full --clock-control none --import-source on -k 'regex:pair_kernel|stream_kernel|maxcurve_edge_kernel' -s 1 -c 1 -f -o gpurun_out/${TAG}_ncu_$wl \
      python bench.py --workload $wl --steps 2 --warmup 1 --no-cpu-baseline --no-configs --sustain 0 --e2e-frames 8 --e2e-steps 1 > gpurun_out/${TAG}_ncu_$wl.log 2>&1
  ncu -i gpurun_out/${TAG}_ncu_$wl.ncu-rep --page raw --csv > gpurun_out/${TAG}_ncu_${wl}_raw.csv 2>/dev/null
  ncu -i gpurun_out/${TAG}_ncu_$wl.ncu-rep --page source --csv > gpurun_out/${TAG}_ncu_${wl}_source.csv 2>/dev/null
  rm -f gpurun_out/${TAG}_ncu_$wl.ncu-rep
done
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
tail -c 600 gpurun_out/${TAG}_bench_ref.json
ls -la gpurun_out | tail -50
