#!/bin/bash
# Round 2, GPU batch 7: the 16-bit kernels inside the +-1 LSB contract (flexion with
# MUFU.RSQ normalisation, bearing / max-curve quotient without the reciprocal's Newton
# step): whole GPU suite, then A/B against the bit-exact forms.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -x 2>&1 | tail -30 > gpurun_out/r2b7_pytest.txt
tail -6 gpurun_out/r2b7_pytest.txt
export BENCH_ARGS="--e2e-frames 8 --e2e-steps 1"
LIBS="base refine" WLS="bearing4_960x540 maxcurve_960x540 range_maxcurve_4096x2160" bash tools/ab_gpu.sh r2b7
LIBS="base flex2" WLS="flexion_960x540" bash tools/ab_gpu.sh r2b7f
DC_EXACT_U16=1 LIBS="base" WLS="flexion_960x540" bash tools/ab_gpu.sh r2b7x
python bench.py --workload bearing4_f64_960x540 --steps 5 --warmup 3 --no-cpu-baseline --no-configs --sustain 0 --e2e-frames 8 --e2e-steps 1 2>&1 | cut -c1-600
python bench.py --workload flexion_f64_960x540 --steps 5 --warmup 3 --no-cpu-baseline --no-configs --sustain 0 --e2e-frames 8 --e2e-steps 1 2>&1 | cut -c1-600
