#!/bin/bash
mkdir -p gpurun_out
python tools/sim_rank.py --world 2 --steps 20 > gpurun_out/sim_w2.json 2> gpurun_out/sim_w2.err; tail -2 gpurun_out/sim_w2.err
python tools/sim_rank.py --world 8 --steps 20 > gpurun_out/sim_w8.json 2> gpurun_out/sim_w8.err
python tools/sim_rank.py --world 8 --steps 160 > gpurun_out/sim_w8_160.json 2> gpurun_out/sim_w8_160.err
python - <<'PY'
import json
for f in ("sim_w2","sim_w8","sim_w8_160"):
    d=json.load(open("gpurun_out/%s.json"%f))
    for k,v in d.items():
        if isinstance(v,dict): print(f,k,{a:round(b,4) for a,b in v.items()})
PY
