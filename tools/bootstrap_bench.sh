#!/bin/bash
# The reference's CKKS EvalBootstrap (parameters of src/pke/examples/simple-ckks-bootstrapping.cpp, full packing) at larger rings,
# CPU path with all host threads against every DCRTPoly on the GPU.  -> gpurun_out/${1}_bootstrap.jsonl
R=${1:-r02y}; O=gpurun_out; T=$(nproc); mkdir -p $O; : > $O/${R}_bootstrap.jsonl
for n in ${2:-16384 32768 65536}; do
  for att in 1 0; do
    FHE_B200_ATTACH=$att OMP_NUM_THREADS=$T timeout 1500 ./oracle/_ref/pke-driver-hooked bootstrap $n $((n/2)) 2 2>> $O/${R}_bootstrap.err | tee -a $O/${R}_bootstrap.jsonl | cut -c1-300
  done
done
grep digest $O/${R}_bootstrap.err | tail -12
