python tools/prof_ntt.py 16 30 64 1 3 > gpurun_out/r02c_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ws_kernel -s 8 -c 4 -o gpurun_out/r02c_prof_ws python tools/prof_ntt.py 16 30 64 1 3 > gpurun_out/r02c_ncu.log 2>&1
tail -3 gpurun_out/r02c_ncu.log
