#!/bin/bash
# Round 2, GPU batch 6: the no-alloc entry point's test alone in a fresh process,
# then the whole GPU suite (it runs last), memcheck / racecheck of the new kernels,
# the operand-form microbenchmark again (raw output for profiles/), default bench line.
set -u
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short 2>&1 | tail -30 > gpurun_out/r2b6_noalloc_alone.txt
tail -5 gpurun_out/r2b6_noalloc_alone.txt
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -60 > gpurun_out/r2b6_pytest.txt
tail -4 gpurun_out/r2b6_pytest.txt
for tool in memcheck racecheck; do
  timeout 600 compute-sanitizer --tool $tool python -m pytest tests/test_zz_no_alloc.py -m gpu -q -k "two_calls or errors" 2>&1 | tail -12 > gpurun_out/r2b6_sanitizer_$tool.txt
  tail -4 gpurun_out/r2b6_sanitizer_$tool.txt
done
tests/cuda/bin/pipebench > gpurun_out/r2_pipebench_rf_b200.txt 2>&1
tail -14 gpurun_out/r2_pipebench_rf_b200.txt
python bench.py > gpurun_out/r2b6_bench_default.json 2> gpurun_out/r2b6_bench_default.err
tail -c 2500 gpurun_out/r2b6_bench_default.json; tail -3 gpurun_out/r2b6_bench_default.err
