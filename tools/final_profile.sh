#!/bin/bash
# Round-end measurement: bench lines for every workload, default line with CPU
# baseline, reference arm, ncu launch list and one --set full capture per kernel.
set -u
mkdir -p gpurun_out
TAG=${1:-r1f}
python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/${TAG}_pytest_gpu.txt
for wl in bearing4_960x540 flexion_960x540 curvature_1920x1080 range_maxcurve_4096x2160 range_960x540 maxcurve_960x540 all_1920x1080; do
  python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err
done
python bench.py > gpurun_out/${TAG}_bench_default.json 2> gpurun_out/${TAG}_bench_default.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bearing4.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launches.log 2>&1
for wl in bearing4_960x540 flexion_960x540 curvature_1920x1080 range_maxcurve_4096x2160 maxcurve_960x540 range_960x540; do
  ncu --set full --clock-control none --import-source on -k 'regex:pair_kernel|stream_kernel|maxcurve_edge_kernel' -s 1 -c 1 -f -o gpurun_out/${TAG}_ncu_$wl \
      python bench.py --workload $wl --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_ncu_$wl.log 2>&1
  # gpurun_out travels back only below 64 MiB: keep the CSV pages, drop the report
  ncu -i gpurun_out/${TAG}_ncu_$wl.ncu-rep --page raw --csv > gpurun_out/${TAG}_ncu_${wl}_raw.csv 2>/dev/null
  ncu -i gpurun_out/${TAG}_ncu_$wl.ncu-rep --page source --csv > gpurun_out/${TAG}_ncu_${wl}_source.csv 2>/dev/null
  if [ $wl != bearing4_960x540 ]; then rm -f gpurun_out/${TAG}_ncu_$wl.ncu-rep; fi
done
ls -la gpurun_out | tail -30
