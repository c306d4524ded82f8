#!/bin/bash
# quick look: attached scheme bench at two small configurations + the tests closest to the allocator
R=${1:-r03g}; O=gpurun_out; T=$(nproc); mkdir -p $O; : > $O/${R}_pke_quick.jsonl
timeout 900 python -m pytest tests/test_gpu_keyswitch.py tests/test_live_intercept.py tests/test_pke_under_hook.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3
for cfg in "bfv-hps 16384 3" "ckks 16384 6" "ckks 65536 24"; do
  FHE_B200_ATTACH=1 OMP_NUM_THREADS=$T timeout 600 ./oracle/_ref/pke-driver-hooked bench $cfg $T 64 2>> $O/${R}_pke_quick.err | tee -a $O/${R}_pke_quick.jsonl | cut -c1-200
done
