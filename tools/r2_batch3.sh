#!/bin/bash
# Round 2, GPU batch 3: tests on the +0-addend / single-exit kernels, operand-form
# microbenchmark, A/B against the round-1 kernels, depth2x start-up profile, full bench line.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -60 > gpurun_out/r2b3_pytest.txt
tail -4 gpurun_out/r2b3_pytest.txt
tests/cuda/bin/pipebench > gpurun_out/r2_pipebench_rf_b200.txt 2>&1
tail -14 gpurun_out/r2_pipebench_rf_b200.txt
export BENCH_ARGS="--e2e-frames 8 --e2e-steps 2"
LIBS="base cur multi negz" WLS="bearing4_960x540 flexion_960x540 curvature_1920x1080 maxcurve_960x540 range_maxcurve_4096x2160" bash tools/ab_gpu.sh r2b3
LIBS="mcedouble" WLS="maxcurve_960x540 range_maxcurve_4096x2160" bash tools/ab_gpu.sh r2b3m
LIBS="flex4" WLS="flexion_960x540" bash tools/ab_gpu.sh r2b3f
LIBS="cur" WLS="bearing4_f32_960x540 flexion_f32_960x540" bash tools/ab_gpu.sh r2b3l
# depth2x start-up (BASELINE config 1: 100 frames 960x540 flexion), new order vs context first
python tools/bench_depth2x.py --frames 100 --size 960x540 --features flexion --workdir /dev/shm/c1 > gpurun_out/r2b3_depth2x_c1.jsonl 2>&1
cd /dev/shm/c1
for i in 1 2 3; do /usr/bin/time -f "wall %e s" env DEPTH2X_PROFILE=1 $OLDPWD/depth-conversions_b200/bin/depth2x -c intrinsic.txt -i data{}-depth.png -s 0 -e 99 flexion -o out/f-{}.png; done > $OLDPWD/gpurun_out/r2b3_depth2x_profile_overlap.txt 2>&1
for i in 1 2 3; do /usr/bin/time -f "wall %e s" env DEPTH2X_PROFILE=1 DEPTH2X_SERIAL_INIT=1 $OLDPWD/depth-conversions_b200/bin/depth2x -c intrinsic.txt -i data{}-depth.png -s 0 -e 99 flexion -o out/f-{}.png; done > $OLDPWD/gpurun_out/r2b3_depth2x_profile_serial.txt 2>&1
cd $OLDPWD
cat gpurun_out/r2b3_depth2x_c1.jsonl | cut -c1-400
grep -E "wall|dc_create|pipeline drained|run_batch end|^\[profile.*main" gpurun_out/r2b3_depth2x_profile_overlap.txt | head -30
python bench.py > gpurun_out/r2b3_bench_default.json 2> gpurun_out/r2b3_bench_default.err
tail -c 3000 gpurun_out/r2b3_bench_default.json; tail -5 gpurun_out/r2b3_bench_default.err
