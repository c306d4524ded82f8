#!/bin/bash
# tensor-core basis switch: parity first (bounded), then timing against the integer kernel
R=${1:-r02k}; O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_basis.py -m gpu -x -q > $O/${R}_pytest_basis_tc.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_basis_tc.log
tail -15 $O/${R}_pytest_basis_tc.log
if grep -q "rc=0" $O/${R}_pytest_basis_tc.log; then
  timeout 600 python -m pytest tests/test_gpu_poverq.py tests/test_gpu_keyswitch.py tests/test_live_intercept.py -m gpu -x -q > $O/${R}_pytest_more_tc.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_more_tc.log
  tail -4 $O/${R}_pytest_more_tc.log
  FHE_B200_BASIS_TC=0 timeout 300 python tools/prof_basis.py 16 30 10 16 5 > $O/${R}_basis_int.json 2>&1
  FHE_B200_BASIS_TC=1 timeout 300 python tools/prof_basis.py 16 30 10 16 5 > $O/${R}_basis_tc.json 2>&1
  cat $O/${R}_basis_int.json; echo; cat $O/${R}_basis_tc.json
fi
