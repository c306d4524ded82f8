#!/bin/bash
# The reference's own examples (sources untouched, oracle/_ref/example-*) with every DCRTPoly on the GPU against the CPU path:
# same printed output (deterministic PRNG, one thread), wall-clock of each.
for ex in simple-ckks-bootstrapping simple-integers; do
  B=./oracle/_ref/example-$ex
  [ -x $B ] || { echo "$ex: not built"; continue; }
  s=$(date +%s.%N); OMP_NUM_THREADS=1 FHE_B200_ATTACH=0 timeout 900 $B > gpurun_out/ex_${ex}_cpu.txt 2> gpurun_out/ex_${ex}_cpu.err; rc0=$?; e=$(date +%s.%N)
  t0=$(python -c "print(round($e - $s, 2))")
  s=$(date +%s.%N); OMP_NUM_THREADS=1 FHE_B200_ATTACH=1 FHE_B200_ATTACH_REPORT=1 timeout 900 $B > gpurun_out/ex_${ex}_gpu.txt 2> gpurun_out/ex_${ex}_gpu.err; rc1=$?; e=$(date +%s.%N)
  t1=$(python -c "print(round($e - $s, 2))")
  if cmp -s gpurun_out/ex_${ex}_cpu.txt gpurun_out/ex_${ex}_gpu.txt; then same=identical; else same=DIFFERENT; fi
  echo "$ex: cpu rc=$rc0 ${t0}s, attached rc=$rc1 ${t1}s, outputs $same"
  tail -3 gpurun_out/ex_${ex}_gpu.txt; tail -2 gpurun_out/ex_${ex}_gpu.err
done
