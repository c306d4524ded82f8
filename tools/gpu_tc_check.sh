#!/bin/bash
# tensor-core basis switch: parity first (bounded), then timing against the integer kernel
R=${1:-r02k}; O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_basis.py -m gpu -x -q > $O/${R}_pytest_basis_tc.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_basis_tc.log
tail -15 $O/${R}_pytest_basis_tc.log
if grep -q "rc=0" $O/${R}_pytest_basis_tc.log; then
  timeout 600 python -m pytest tests/test_gpu_poverq.py tests/test_gpu_keyswitch.py tests/test_live_intercept.py -m gpu -x -q > $O/${R}_pytest_more_tc.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_more_tc.log
  tail -4 $O/${R}_pytest_more_tc.log
  FHE_B200_BASIS_TC=0 timeout 300 python tools/prof_basis.py 16 30 10 16 5 > $O/${R}_basis_int.json 2>&1
  for tpr in 1 2 3; do FHE_B200_BASIS_TC_SUB=$tpr timeout 300 python tools/prof_basis.py 16 30 10 16 5 > $O/${R}_basis_tc_tpr$tpr.json 2>&1; done
  python - <<PY
import json
for f in ["int","tc_tpr1","tc_tpr2","tc_tpr3"]:
    d=json.load(open("$O/${R}_basis_%s.json"%f))
    print(f, {k:round(v["ms"],4) for k,v in d.items() if isinstance(v,dict)})
PY
fi
