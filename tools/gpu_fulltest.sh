#!/bin/bash
R=${1:-r02h}; O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${R}_pytest_gpu.log 2>&1; echo "rc=$?" >> $O/${R}_pytest_gpu.log
tail -8 $O/${R}_pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > $O/${R}_bench.json 2> $O/${R}_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/'+"R".replace("R","${R}")+'_bench.json')) if False else None
PY
python -c "
import json,sys
d=json.load(open('$O/${R}_bench.json'))
print(d['value'], d['e2e']['value'], d['roofline']['frac'])
print(json.dumps(d['extra'].get('keyswitch_hybrid_n65536_ql30_p10_alpha10_b16')), d['extra'].get('error'))
"
