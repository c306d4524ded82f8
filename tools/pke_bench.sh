#!/bin/bash
# The reference's own EvalAdd / EvalMult+relinearisation / Rescale / EvalRotate over independent ciphertexts, CPU path against
# every DCRTPoly on the GPU (oracle/_ref/pke-driver-hooked bench ...).  Output: JSON lines -> gpurun_out/${1}_pke_bench.jsonl
R=${1:-r02u}; O=gpurun_out; T=$(nproc); mkdir -p $O; : > $O/${R}_pke_bench.jsonl
for cfg in "bfv-bv 16384 3" "bfv-bv 32768 8" "bfv-hps 16384 3" "bfv-behz 16384 3" "bfv-hpspoverq 16384 3" "ckks 16384 6" "bfv-hps 32768 8" "ckks 32768 12" "ckks 65536 24"; do
  for att in 0 1; do
    FHE_B200_ATTACH=$att OMP_NUM_THREADS=$T timeout 900 ./oracle/_ref/pke-driver-hooked bench $cfg $T 64 2>> $O/${R}_pke_bench.err | tee -a $O/${R}_pke_bench.jsonl | cut -c1-230
  done
done
