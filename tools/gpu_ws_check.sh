#!/bin/bash
# GPU check of the persistent NTT kernels: parity (default dispatch, then forced), timing against the one-tile kernels.
R=${1:-r02b}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "ntt" > $O/${R}_pytest_ntt.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_ntt.log
tail -3 $O/${R}_pytest_ntt.log
for cfg in "16 30 1024" "16 30 64" "15 16 128" "14 8 256"; do
  for algo in 2 3; do
    timeout 300 python tools/prof_ntt.py $cfg 5 $algo >> $O/${R}_ntt_timing.log 2>&1
  done
  FHE_B200_NTT_WS_GROUPS=5 timeout 300 python tools/prof_ntt.py $cfg 5 3 >> $O/${R}_ntt_timing.log 2>&1
done
cat $O/${R}_ntt_timing.log
FHE_B200_NTT_WS=1 timeout 900 python -m pytest tests -m gpu -x -q > $O/${R}_pytest_gpu_ws.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_gpu_ws.log
tail -3 $O/${R}_pytest_gpu_ws.log
tools/bin/int_probe > $O/${R}_int_probe.jsonl 2>&1
cat $O/${R}_int_probe.jsonl
