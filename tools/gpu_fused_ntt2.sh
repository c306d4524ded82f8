#!/bin/bash
R=${1:-r03r}; O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/${R}_pytest_gpu.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_gpu.log; tail -3 $O/${R}_pytest_gpu.log
FHE_B200_NTT_FUSED=2 timeout 300 ncu --set full --clock-control none -k regex:ntt_fused -s 4 -c 2 -o $O/${R}_prof_ntt_fused python tools/prof_ntt.py 14 8 1024 1 > $O/${R}_ncu_fused.log 2>&1
ncu -i $O/${R}_prof_ntt_fused.ncu-rep --page raw --csv > $O/${R}_ntt_fused_n16384_t8_b1024_full.csv 2>/dev/null
rm -f $O/${R}_prof_ntt_fused.ncu-rep
python - <<PY
import csv
rows=list(csv.reader(open('$O/${R}_ntt_fused_n16384_t8_b1024_full.csv')))
h=rows[0]
want=['Kernel Name','dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active']
idx=[h.index(w) for w in want if w in h]
for r in rows[2:]:
    print([r[i][:40] for i in idx])
print([h[i] for i in idx]); print(rows[1][idx[1]] if len(idx)>1 else '')
PY
