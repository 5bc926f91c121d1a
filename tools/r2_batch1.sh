#!/bin/bash
# Round 2, GPU batch 1: operand-bandwidth microbenchmark, ablation timings of the
# bearing / flexion kernels, structural variants of the shared-edge max-curve kernel.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2b1_smi.txt
tests/cuda/bin/pipebench > gpurun_out/r2_pipebench_rf_b200.txt 2>&1
tail -25 gpurun_out/r2_pipebench_rf_b200.txt
export BENCH_ARGS="--e2e-frames 8"
LIBS="base abl1 abl2 abl4 abl6 abl7 approx approx7 carrysq" WLS="bearing4_960x540" bash tools/ab_gpu.sh r2b1_bearing
LIBS="base abl1" WLS="flexion_960x540 curvature_1920x1080" bash tools/ab_gpu.sh r2b1_flex
LIBS="base mcesmem3 mcesingle mcesmem3single mcesmem2" WLS="maxcurve_960x540 range_maxcurve_4096x2160" bash tools/ab_gpu.sh r2b1_mce
