#!/bin/bash
R=${1:-r02g}; O=gpurun_out
run() { echo "== $*" >> $O/${R}_race.log; env "$@" python tools/hook_bfv_bench.py 64 hps 16 2>>$O/${R}_race.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bit_exact', d['bit_exact'], 'b200/s', round(d['b200']['ciphertext_products_per_s']), 'speedup', round(d['speedup'],1))" >> $O/${R}_race.log 2>&1; }
for i in 1 2 3; do run X=1; done
run FHE_B200_HOOK_SYNC=1
run FHE_B200_HOOK_ONE_STREAM=1
run FHE_B200_HOOK_SYNC=1 FHE_B200_HOOK_ONE_STREAM=1
for i in 1 2; do echo "== hpspoverq" >> $O/${R}_race.log; python tools/hook_bfv_bench.py 64 hpspoverq 16 2>>$O/${R}_race.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('bit_exact', d['bit_exact'], round(d['speedup'],1))" >> $O/${R}_race.log; done
cat $O/${R}_race.log; tail -n 5 $O/${R}_race.err
