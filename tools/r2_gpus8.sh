#!/bin/bash
# Round 2, the ONE 8-GPU call: host<->device copy ceiling
# of the box with 1/2/4/8 GPUs active, the multi-device depth2x test un-skipped,
# BASELINE config 5 through depth2x --devices 8/4/2/1, bench.py e2e at N = 1/2/4/8.
set -u
FRAMES=${1:-1000}
DEVS=${DEVS:-8,4,2,1}      # depth2x --devices values, largest first
MAXG=${MAXG:-8}            # GPUs the copy probe uses
BENCH_NS=${BENCH_NS:-}     # e.g. "1 2 4 8": bench.py e2e per N (the driver's SCALE run does this anyway)
mkdir -p gpurun_out
{ nproc; free -g | head -2; df -h /dev/shm | tail -1; lscpu | grep -E "Model name|NUMA node|Socket"; } > gpurun_out/r2c5_box.txt 2>&1
nvidia-smi topo -m >> gpurun_out/r2c5_box.txt 2>&1
cat gpurun_out/r2c5_box.txt | head -12
tests/cuda/bin/pcie_probe $MAXG 64 20 > gpurun_out/r2c5_pcie_probe.txt 2>&1
cat gpurun_out/r2c5_pcie_probe.txt
python -m pytest tests/test_depth2x_gpu.py -m gpu -q --tb=short -k devices 2>&1 | tail -8 > gpurun_out/r2c5_pytest_depth2x.txt
tail -3 gpurun_out/r2c5_pytest_depth2x.txt
python tools/config5.py --frames $FRAMES --devices $DEVS --out gpurun_out/r2c5_depth2x.jsonl 2>&1 | cut -c1-330
for n in $BENCH_NS; do
  if [ $n = 1 ]; then
    python bench.py --gpus 1 --steps 20 --warmup 3 --no-configs --sustain 0 --no-cpu-baseline > gpurun_out/r2c5_bench_n$n.json 2> gpurun_out/r2c5_bench_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 3 --no-configs --sustain 0 --no-cpu-baseline > gpurun_out/r2c5_bench_n$n.json 2> gpurun_out/r2c5_bench_n$n.err
  fi
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r2c5_bench_n$n.json").read().strip().splitlines()[-1])
    print("N=$n value", round(d["value"]), "MPix/s ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), d["clocks"]["reasons"])
except Exception as e:
    print("N=$n FAILED", e); print(open("gpurun_out/r2c5_bench_n$n.err").read()[-1500:])
PY
done
