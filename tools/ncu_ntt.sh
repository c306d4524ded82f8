#!/bin/bash
# One ncu --set full capture of the four NTT kernels (forward + inverse) at N=2^16, T=30, B=${1:-64}; report -> gpurun_out/
B=${1:-64}; R=${2:-r02t}; ALGO=${3:-0}
python tools/prof_ntt.py 16 30 $B 1 $ALGO > gpurun_out/${R}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ntt_ -s 8 -c 4 -o gpurun_out/${R}_ntt python tools/prof_ntt.py 16 30 $B 1 $ALGO > gpurun_out/${R}_ncu.log 2>&1
tail -3 gpurun_out/${R}_plain.log; tail -3 gpurun_out/${R}_ncu.log
