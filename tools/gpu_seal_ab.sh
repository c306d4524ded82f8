#!/bin/bash
# one host thread through the live intercept: how the arena's seal group and growth limit change the BFV product rate
O=gpurun_out; R=${1:-r03m}; : > $O/${R}_seal_ab.txt
for cfg in "0 46000" "1 46000" "32 46000" "1 0" "0 0"; do
  set -- $cfg
  FHE_B200_ARENA_SEAL=$1 FHE_B200_ARENA_GROW_MB=$2 timeout 300 python tools/hook_bfv_bench.py 64 hps 1 2>> $O/${R}_seal_ab.err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('seal $1 grow_mb $2 threads', d['b200']['threads'], 'cpu', round(d['cpu']['ciphertext_products_per_s'],1), 'b200', round(d['b200']['ciphertext_products_per_s'],1), d['bit_exact'])" | tee -a $O/${R}_seal_ab.txt
done
FHE_B200_ARENA_SEAL=0 timeout 300 python tools/hook_bfv_bench.py 64 hps 16 2>> $O/${R}_seal_ab.err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('adaptive, 16 threads: b200', round(d['b200']['ciphertext_products_per_s'],1), d['bit_exact'])" | tee -a $O/${R}_seal_ab.txt
