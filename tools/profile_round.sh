#!/bin/bash
# Round profile pass, run on the GPU box:  gpurun -- 'bash tools/profile_round.sh r02'
# Produces under gpurun_out/: the integer-pipe probe, the GPU test log, the bench line of both arms, the ncu launch
# list of the same bench command, full ncu captures of the NTT and element-wise kernels at the bench shape and of the
# tensor-core basis switch, and the live-intercept benches.  Summarise here with tools/ncu_summary.py and copy what
# should be judged into profiles/.
R=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/${R}_gpu.txt 2>&1
nproc > $O/${R}_host.txt; free -g >> $O/${R}_host.txt
[ -x tools/bin/int_probe ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/int_probe tools/int_probe.cu
tools/bin/int_probe > $O/${R}_int_probe.jsonl 2>&1
python -m pytest tests -m gpu -x -q > $O/${R}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${R}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${R}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${R}_smoke.log
python bench.py --impl reference --steps 20 --warmup 5 > $O/${R}_bench_ref.json 2> $O/${R}_bench_ref.err
python bench.py --steps 20 --warmup 5 > $O/${R}_bench.json 2> $O/${R}_bench.err
python bench.py --steps 2 --warmup 3 --no-extra > $O/${R}_bench_short.json 2> $O/${R}_bench_short.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/${R}_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extra > $O/${R}_ncu_launch.log 2>&1
python tools/prof_ntt.py 16 30 1024 1 > $O/${R}_plain_ntt.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ntt_ -s 8 -c 4 -o $O/${R}_prof_ntt \
    python tools/prof_ntt.py 16 30 1024 1 > $O/${R}_ncu_ntt.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ew_ring -s 6 -c 2 -o $O/${R}_prof_ew \
    python bench.py --steps 2 --warmup 3 --no-extra > $O/${R}_ncu_ew.log 2>&1
python tools/prof_basis.py 16 30 10 16 2 > $O/${R}_plain_basis.json 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'basis_switch|ks_inner|decrypt_scale' -c 6 \
    -o $O/${R}_prof_basis python tools/prof_basis.py 16 30 10 16 1 > $O/${R}_ncu_basis.log 2>&1
FHE_B200_BASIS_TC=0 python tools/prof_basis.py 16 30 10 16 5 > $O/${R}_basis_int.json 2>&1
python tools/prof_basis.py 16 30 10 16 5 > $O/${R}_basis_tc.json 2>&1
for t in 1 16; do
  python tools/hook_bench.py 14 8 64 4 $t >> $O/${R}_hook_bench.jsonl 2>> $O/${R}_hook.err
  python tools/hook_bench.py 16 30 16 4 $t >> $O/${R}_hook_bench.jsonl 2>> $O/${R}_hook.err
  python tools/hook_bfv_bench.py 64 hps $t >> $O/${R}_hook_bfv.jsonl 2>> $O/${R}_hook.err
  python tools/hook_bfv_bench.py 64 hpspoverq $t >> $O/${R}_hook_bfv.jsonl 2>> $O/${R}_hook.err
done
bash tools/polybench.sh > $O/${R}_polybench.log 2>&1
ls -la $O | tail -30
