#!/bin/bash
mkdir -p gpurun_out
(timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --workload c4 --steps 20 --warmup 3 --no-c5 > gpurun_out/r02_n2_c4.json 2> gpurun_out/r02_n2_c4.err; echo "c4 rc=$?"); tail -2 gpurun_out/r02_n2_c4.err
(timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --workload c4 --steps 20 --warmup 3 --no-c5 --partition bands --submission sequence > gpurun_out/r02_n2_c4_bands.json 2> gpurun_out/r02_n2_c4_bands.err; echo "c4 bands rc=$?"); tail -2 gpurun_out/r02_n2_c4_bands.err
python - <<'PY'
import json
for n in ("r02_n2_c4","r02_n2_c4_bands"):
    d=json.load(open("gpurun_out/%s.json"%n))
    print(n, d["config"]["partition"], "headline %.4f (%s) seq %.4f frames %.4f nccl %.4f e2e fps %.0f"%(d["ms_per_step"], d["config"]["submission_choice"], d["sequence_api"]["ms_per_step"], d["per_frame_api"]["ms_per_step"], d["nccl_gather_baseline"]["ms_per_step"], d["e2e"]["fps"]), {k:v for k,v in d["verify"].items() if k!="what"})
PY
