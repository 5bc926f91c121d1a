#!/bin/bash
# Round 2, GPU batch 2: the whole GPU test suite on the new code, then batch 1.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short --deselect tests/test_ref_shell_scripts.py 2>&1 | tail -150 > gpurun_out/r2b2_pytest.txt
python -m pytest tests/test_ref_shell_scripts.py -m gpu -q --tb=short 2>&1 | tail -100 > gpurun_out/r2b2_pytest_shell.txt
tail -5 gpurun_out/r2b2_pytest.txt; tail -5 gpurun_out/r2b2_pytest_shell.txt
bash tools/r2_batch1.sh
