#!/bin/bash
# A/B of attached bootstrapping at one ring: key-switch graph replay off/on, arena growth off/on.
R=${1:-r02x}; N=${2:-16384}; O=gpurun_out; T=$(nproc); mkdir -p $O; : > $O/${R}_boot_ab.jsonl
for cfg in "0 0" "0 46000" "1 0" "1 46000" "2 46000"; do
  set -- $cfg
  FHE_B200_KS_GRAPH=$1 FHE_B200_ARENA_GROW_MB=$2 FHE_B200_ATTACH=1 OMP_NUM_THREADS=$T timeout 600 ./oracle/_ref/pke-driver-hooked bootstrap $N $((N/2)) 3 2>> $O/${R}_boot_ab.err \
    | grep "^{" | sed "s/^{/{\"ks_graph\": $1, \"arena_grow_mb\": $2, /" | tee -a $O/${R}_boot_ab.jsonl | cut -c1-260
done
