#!/bin/bash
# L2 chunking sweep of the two-pass NTT at the bench shape (N=2^16, T=30, B=1024) and B=64.
O=gpurun_out; R=${1:-r02d}
for mb in 0 16 32 48 64 96; do
  echo "L2_MB=$mb" >> $O/${R}_l2.log
  FHE_B200_NTT_L2_MB=$mb timeout 300 python tools/prof_ntt.py 16 30 1024 5 2 >> $O/${R}_l2.log 2>&1
  FHE_B200_NTT_L2_MB=$mb timeout 300 python tools/prof_ntt.py 15 16 512 5 2 >> $O/${R}_l2.log 2>&1
  FHE_B200_NTT_L2_MB=$mb timeout 300 python tools/prof_ntt.py 14 8 1024 5 2 >> $O/${R}_l2.log 2>&1
done
cat $O/${R}_l2.log
