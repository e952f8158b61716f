#!/bin/bash
# eight-GPU lease: the default bench line at N = 8 (C3 headline + C5 record)
mkdir -p gpurun_out; nvidia-smi -L | wc -l
(NCCL_DEBUG=INFO NCCL_DEBUG_FILE=gpurun_out/r02_n8_nccl_%h_%p.log timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r02_n8_bench.json 2> gpurun_out/r02_n8_bench.err; echo "bench rc=$?"); tail -3 gpurun_out/r02_n8_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_n8_bench.json"))
print("n8 seq ms/frame %.4f"%d["ms_per_step"], "value %.1f"%d["value"], "repeats", d["repeats_ms_per_step"]["median"], "per-frame-api %.4f"%d["per_frame_api"]["ms_per_step"], "nccl %.4f"%d["nccl_gather_baseline"]["ms_per_step"], "e2e fps %.0f"%d["e2e"]["fps"])
print(" verify", {k:v for k,v in d["verify"].items() if k!="what"}); print(" kernel_vs_ideal", d["kernel_vs_ideal"]); print(" c5", {k:v for k,v in d.get("c5",{}).items() if k in ("ms_per_step","gsamples","verify")})
PY
rm -f gpurun_out/r02_n8_nccl_*   # keep the merge small
