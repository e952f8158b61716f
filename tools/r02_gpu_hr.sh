#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_hostring.py tests/test_bands.py -m gpu -q -rs 2>&1 | tail -3
(timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 3 --no-c5 > gpurun_out/r02_n2_hr.json 2> gpurun_out/r02_n2_hr.err; echo "bench rc=$?"); tail -2 gpurun_out/r02_n2_hr.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_n2_hr.json"))
print("headline %.4f e2e fps %.0f (device fps %.0f)"%(d["ms_per_step"], d["e2e"]["fps"], d["fps"]), {k:v for k,v in d["verify"].items() if k!="what"})
PY
