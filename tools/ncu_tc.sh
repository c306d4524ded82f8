python tools/prof_basis.py 16 30 10 16 2 > gpurun_out/r02o_plain.json 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:basis_switch_tc -s 4 -c 2 -o gpurun_out/r02o_prof_tc python tools/prof_basis.py 16 30 10 16 2 > gpurun_out/r02o_ncu.log 2>&1
tail -3 gpurun_out/r02o_ncu.log; python -c "
import json; d=json.load(open('gpurun_out/r02o_plain.json')); print({k:round(v['ms'],4) for k,v in d.items() if isinstance(v,dict)})"
