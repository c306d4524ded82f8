#!/bin/bash
# The captured key-switch middle (csrc/runtime.cu keyswitch_graphed): parity tests, then the scheme-level bench with the capture
# off and on.  Output under gpurun_out/${1}_*.
R=${1:-r02v}; O=gpurun_out; T=$(nproc); mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_keyswitch.py tests/test_gpu_basis.py tests/test_live_intercept.py tests/test_pke_under_hook.py -x -q -m gpu > $O/${R}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 $O/${R}_pytest.log
: > $O/${R}_pke_graph_ab.jsonl
for cfg in "bfv-hps 16384 3" "ckks 16384 6" "ckks 32768 12" "ckks 65536 24"; do
  for g in 0 1; do
    FHE_B200_KS_GRAPH=$g FHE_B200_ATTACH=1 OMP_NUM_THREADS=$T timeout 600 ./oracle/_ref/pke-driver-hooked bench $cfg $T 64 2>> $O/${R}_pke_graph_ab.err \
      | sed "s/^{/{\"ks_graph\": $g, /" | tee -a $O/${R}_pke_graph_ab.jsonl | cut -c1-200
  done
done
grep "digest" $O/${R}_pke_graph_ab.err | sort | uniq -c
