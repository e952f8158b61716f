#!/bin/bash
mkdir -p gpurun_out
(timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"); tail -1 gpurun_out/r02_smoke.log
(timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r02_n4_bench.json 2> gpurun_out/r02_n4_bench.err; echo "bench rc=$?"); tail -2 gpurun_out/r02_n4_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_n4_bench.json"))
print("n4 headline %.4f (%s) value %.1f seq %.4f repeats %.4f frames %.4f nccl %.4f e2e fps %.0f"%(d["ms_per_step"], d["config"]["submission_choice"], d["value"], d["sequence_api"]["ms_per_step"], d["repeats_ms_per_step"]["median"], d["per_frame_api"]["ms_per_step"], d["nccl_gather_baseline"]["ms_per_step"], d["e2e"]["fps"]))
print({k:v for k,v in d["verify"].items() if k!="what"}, d["c5"]["ms_per_step"], d["c5"]["verify"]["ok"])
PY
