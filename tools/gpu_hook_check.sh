#!/bin/bash
R=${1:-r02f}; O=gpurun_out
nproc > $O/${R}_host.txt
timeout 900 python -m pytest tests/test_live_intercept.py -m gpu -x -q > $O/${R}_pytest_hook.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_hook.log
tail -4 $O/${R}_pytest_hook.log
for t in 1 16; do
  timeout 600 python tools/hook_bench.py 14 8 64 4 $t >> $O/${R}_hook_bench.jsonl 2>> $O/${R}_hook_bench.err
  timeout 600 python tools/hook_bench.py 16 30 16 4 $t >> $O/${R}_hook_bench.jsonl 2>> $O/${R}_hook_bench.err
  timeout 600 python tools/hook_bfv_bench.py 64 hps $t >> $O/${R}_hook_bfv.jsonl 2>> $O/${R}_hook_bfv.err
  timeout 600 python tools/hook_bfv_bench.py 64 hpspoverq $t >> $O/${R}_hook_bfv.jsonl 2>> $O/${R}_hook_bfv.err
done
cat $O/${R}_hook_bench.jsonl $O/${R}_hook_bfv.jsonl; tail -5 $O/${R}_hook_bench.err $O/${R}_hook_bfv.err
