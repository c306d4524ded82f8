#!/bin/bash
# Multi-GPU check (gpurun --gpus N): group API + C++ surface over N devices, bench under torchrun and in one process.
N=${1:-2}; R=${2:-r02e}; O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/${R}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_group.py -m gpu -x -q > $O/${R}_pytest_group.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_group.log
tail -5 $O/${R}_pytest_group.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-extra > $O/${R}_bench_torchrun_${N}gpu.json 2> $O/${R}_bench_torchrun_${N}gpu.err; echo "torchrun rc=$?"
tail -c 1500 $O/${R}_bench_torchrun_${N}gpu.json; tail -5 $O/${R}_bench_torchrun_${N}gpu.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/${R}_bench_ref_torchrun_${N}gpu.json 2> $O/${R}_bench_ref_torchrun_${N}gpu.err; echo "ref rc=$?"
cat $O/${R}_bench_ref_torchrun_${N}gpu.json | cut -c1-400
timeout 900 python bench.py --gpus $N --steps 5 --warmup 3 > $O/${R}_bench_group_${N}gpu.json 2> $O/${R}_bench_group_${N}gpu.err; echo "group rc=$?"
cat $O/${R}_bench_group_${N}gpu.json; tail -5 $O/${R}_bench_group_${N}gpu.err
