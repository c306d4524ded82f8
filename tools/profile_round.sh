#!/bin/bash
# Round profile pass, run on the GPU box:  gpurun -- 'bash tools/profile_round.sh r02'
# Produces under gpurun_out/: the integer-pipe probe, the GPU test log, the bench line of both arms, the ncu launch
# list of the same bench command, and full ncu captures of the NTT (at the bench shape) and element-wise kernels.
# Summarise here with tools/ncu_summary.py and copy what should be judged into profiles/.
R=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/${R}_gpu.txt 2>&1
nproc > $O/${R}_host.txt; free -g >> $O/${R}_host.txt
[ -x tools/bin/int_probe ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/int_probe tools/int_probe.cu
tools/bin/int_probe > $O/${R}_int_probe.jsonl 2>&1
python -m pytest tests -m gpu -x -q > $O/${R}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${R}_pytest_gpu.log
python bench.py --steps 20 --warmup 5 > $O/${R}_bench.json 2> $O/${R}_bench.err
python bench.py --impl reference --steps 5 --warmup 1 > $O/${R}_bench_ref.json 2> $O/${R}_bench_ref.err
python bench.py --steps 2 --warmup 3 --no-extra > $O/${R}_bench_short.json 2> $O/${R}_bench_short.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/${R}_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extra > $O/${R}_ncu_launch.log 2>&1
python tools/prof_ntt.py 16 30 1024 1 > $O/${R}_plain_ntt.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ntt_ -s 8 -c 4 -o $O/${R}_prof_ntt \
    python tools/prof_ntt.py 16 30 1024 1 > $O/${R}_ncu_ntt.log 2>&1
ls -la $O | tail -20
