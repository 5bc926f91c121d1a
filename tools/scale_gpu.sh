#!/bin/bash
# bench.py at N = 1, 2, 4, 8 GPUs of one box (as the driver launches it), plus the
# all-features workload at N = 8 and depth2x --devices. usage: scale_gpu.sh TAG [NS]
set -u
TAG=${1:-scale}
NS=${2:-"1 2 4 8"}
mkdir -p gpurun_out
for n in $NS; do
  for wl in bearing4_960x540 all_1920x1080; do
    if [ $n = 1 ]; then
      python bench.py --gpus 1 --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_n${n}_$wl.json 2> gpurun_out/${TAG}_n${n}_$wl.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_n${n}_$wl.json 2> gpurun_out/${TAG}_n${n}_$wl.err
    fi
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_n${n}_$wl.json").read().strip().splitlines()[-1])
    print("N=$n $wl value", round(d["value"]), "MPix/s ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), d["clocks"]["reasons"])
except Exception as e:
    print("N=$n $wl FAILED", e); print(open("gpurun_out/${TAG}_n${n}_$wl.err").read()[-1500:])
PY
  done
done | tee gpurun_out/${TAG}_summary.txt
