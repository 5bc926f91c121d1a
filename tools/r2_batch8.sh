#!/bin/bash
# Round 2, GPU batch 8: conflict-free staging layout of the flexion kernels (A/B against the
# slot layout, u16 LSB / u16 exact / f32), then the whole GPU suite.
set -u
mkdir -p gpurun_out
export BENCH_ARGS="--e2e-frames 8 --e2e-steps 1"
LIBS="base flexslots" WLS="flexion_960x540 flexion_f32_960x540" bash tools/ab_gpu.sh r2b8
DC_EXACT_U16=1 LIBS="base flexslots" WLS="flexion_960x540" bash tools/ab_gpu.sh r2b8x
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -30 > gpurun_out/r2b8_pytest.txt
tail -6 gpurun_out/r2b8_pytest.txt
