#!/usr/bin/env bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 30 --warmup 3 > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo bench rc=$?; tail -3 gpurun_out/bench_q.err
python - <<'PY'
import json
b=json.loads(open('gpurun_out/bench_q.json').read().strip().splitlines()[-1])
print(b['value'], b['e2e']['value'])
for k,v in b['extra']['sweeps'].items():
    if isinstance(v,dict): print(k,{a:(round(x,3) if isinstance(x,float) else x) for a,x in v.items()})
print(b['extra']['fast']['sample_direction_sun_and_sky_Gsamples_s'])
PY
