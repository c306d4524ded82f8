#!/bin/bash
# Round profile pass, run on the GPU box:  gpurun -- 'bash tools/profile_round.sh r01'
# Produces under gpurun_out/: the bench line, the ncu launch list of the same bench command, and full ncu
# captures of the element-wise, tensor-product and NTT kernels.  Summarise here with tools/ncu_summary.py.
R=${1:-r01}
O=gpurun_out
mkdir -p $O
tools/bin/int_probe > $O/${R}_int_probe.jsonl 2>&1
python bench.py --steps 20 --warmup 5 > $O/${R}_bench.json 2> $O/${R}_bench.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${R}_bench_launches.csv \
    python bench.py --steps 20 --warmup 5 > $O/${R}_ncu_launch.log 2>&1
python tools/prof_ntt.py 16 30 16 1 > $O/${R}_plain_ntt.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ntt_ -s 8 -c 4 -o $O/${R}_prof_ntt \
    python tools/prof_ntt.py 16 30 16 1 > $O/${R}_ncu_ntt.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'ew_ring|ew_tensor' -s 6 -c 3 -o $O/${R}_prof_ew \
    python bench.py --steps 5 --warmup 3 > $O/${R}_ncu_ew.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ew_tensor -c 1 -o $O/${R}_prof_tensor \
    python bench.py --steps 5 --warmup 3 > $O/${R}_ncu_tensor.log 2>&1
python tools/prof_basis.py 16 30 10 16 2 > $O/${R}_plain_basis.json 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'basis_switch|tower_affine|rescale_spread|automorph' -c 8 \
    -o $O/${R}_prof_basis python tools/prof_basis.py 16 30 10 16 1 > $O/${R}_ncu_basis.log 2>&1
python tools/hook_bench.py 14 8 64 4 > $O/${R}_hook_bench.jsonl 2>&1
python tools/hook_bench.py 16 30 8 4 >> $O/${R}_hook_bench.jsonl 2>&1
ls -la $O | tail -20
