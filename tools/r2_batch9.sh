#!/bin/bash
# Round 2, GPU batch 9: degree-4 acos polynomial of the 16-bit kernels: mathcheck, the parity
# tests, A/B against degree 5.
set -u
mkdir -p gpurun_out
tests/cuda/bin/mathcheck > gpurun_out/r2_mathcheck.txt 2>&1; tail -4 gpurun_out/r2_mathcheck.txt
python -m pytest tests/test_gpu_parity.py tests/test_ref_shell_scripts.py tests/test_depth2x_gpu.py tests/test_gpu_f32_ragged.py tests/test_ref_unittests_gpu.py -m gpu -q --tb=short 2>&1 | tail -15 > gpurun_out/r2b9_pytest.txt
tail -5 gpurun_out/r2b9_pytest.txt
export BENCH_ARGS="--e2e-frames 8 --e2e-steps 1"
LIBS="base deg5" WLS="bearing4_960x540 maxcurve_960x540 range_maxcurve_4096x2160" bash tools/ab_gpu.sh r2b9
