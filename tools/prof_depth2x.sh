#!/bin/bash
# Stage timeline of bin/depth2x (DEPTH2X_PROFILE=1) on synthetic 1080p frames.
set -e; export TIMEFORMAT="wall %R s"
cd /root/repo
N=${1:-400}
python - <<PY
import sys,os
sys.path.insert(0,'tests')
import cv2
from oracle_lib import Oracle,kinect_scaled
o=Oracle()
w,h,n,d=1920,1080,$N,'/dev/shm/p1080'
os.makedirs(d+'/out',exist_ok=True)
k=kinect_scaled(w,h)
open(d+'/intrinsic.txt','w').write(f"{w} {h}\n{k['fx']!r} 0.0 {k['cx']!r}\n0.0 {k['fy']!r} {k['cy']!r}\n0.0 0.0 1.0\n")
fr=[o.synth_depth(0xD2C00001,i,w,h) for i in range(8)]
for i in range(n): cv2.imwrite(f"{d}/data{i}-depth.png",fr[i%8])
PY
cd /dev/shm/p1080
for r in 1 2; do time env DEPTH2X_PROFILE=1 /root/repo/depth-conversions_b200/bin/depth2x -c intrinsic.txt -i data{}-depth.png -s 0 -e $((N-1)) flexion -o out/f-{}.png; done
