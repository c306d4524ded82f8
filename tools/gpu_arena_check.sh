#!/bin/bash
# After the arena's grow-before-stall change: attached bootstrapping at the three rings (attached path only; the CPU figures
# are in profiles/r02_pke_bootstrap.jsonl) and the scheme-level bench at N = 2^16.
R=${1:-r02w}; O=gpurun_out; T=$(nproc); mkdir -p $O; : > $O/${R}_bootstrap.jsonl; : > $O/${R}_pke_bench_n65536.jsonl
for n in 65536 32768 16384; do
  FHE_B200_ATTACH=1 OMP_NUM_THREADS=$T timeout 900 ./oracle/_ref/pke-driver-hooked bootstrap $n $((n/2)) 3 2>> $O/${R}_bootstrap.err | tee -a $O/${R}_bootstrap.jsonl | cut -c1-300
done
grep "digest\|launches" $O/${R}_bootstrap.err | tail -18
for cfg in "ckks 65536 24" "bfv-hps 32768 8"; do
  FHE_B200_ATTACH=1 OMP_NUM_THREADS=$T timeout 600 ./oracle/_ref/pke-driver-hooked bench $cfg $T 64 2>> $O/${R}_pke_bench_n65536.err | tee -a $O/${R}_pke_bench_n65536.jsonl | cut -c1-200
done
