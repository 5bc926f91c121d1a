#!/bin/bash
# BASELINE config 1 end to end: depth2x flexion over 100 PNGs of 960x540,
# three times with the stage profile, beside the reference's CPU path on the same files.
set -u
mkdir -p gpurun_out
python tools/bench_depth2x.py --frames 100 --size 960x540 --features flexion --workdir /dev/shm/c1 > gpurun_out/r2_depth2x_c1.jsonl 2>&1
cut -c1-500 gpurun_out/r2_depth2x_c1.jsonl
R=$PWD
cd /dev/shm/c1 2>/dev/null || { echo "workdir gone"; exit 0; }
ls | head -3
for i in 1 2 3; do TIMEFORMAT="wall %R s"; time env DEPTH2X_PROFILE=1 $R/depth-conversions_b200/bin/depth2x -c intrinsic.txt -i data{}-depth.png -s 0 -e 99 flexion -o out/f-{}.png; done > $R/gpurun_out/r2_depth2x_c1_profile.txt 2>&1
for i in 1 2 3; do TIMEFORMAT="wall %R s"; time env DEPTH2X_PROFILE=1 DEPTH2X_SERIAL_INIT=1 $R/depth-conversions_b200/bin/depth2x -c intrinsic.txt -i data{}-depth.png -s 0 -e 99 flexion -o out/f-{}.png; done > $R/gpurun_out/r2_depth2x_c1_profile_serial.txt 2>&1
cd $R
grep -E "wall|dc_create (begin|end)|pipeline drained|decoded|took" gpurun_out/r2_depth2x_c1_profile.txt | head -40
echo ---- serial init
grep -E "wall|took" gpurun_out/r2_depth2x_c1_profile_serial.txt
