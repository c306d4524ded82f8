#!/bin/bash
# The single-pass NTT kernel (ntt_fused_kernel, N <= 2^14): parity, then time against the two-pass kernels.
R=${1:-r03q}; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_keyswitch.py -x -q -m gpu -k "single_pass or ntt or keyswitch_hybrid" > $O/${R}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${R}_pytest.log
: > $O/${R}_fused_timing.log
for shape in "14 8 1024" "14 8 64" "14 8 4" "14 8 1" "13 8 1024" "13 8 8" "12 8 2048" "12 8 8" "11 4 4096" "10 4 8192"; do
  for f in 0 2; do
    echo -n "fused=$f " >> $O/${R}_fused_timing.log
    FHE_B200_NTT_FUSED=$f timeout 120 python tools/prof_ntt.py $shape 20 0 >> $O/${R}_fused_timing.log 2>&1
  done
done
cat $O/${R}_fused_timing.log
