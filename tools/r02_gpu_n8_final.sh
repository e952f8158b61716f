#!/bin/bash
# eight-GPU lease with the end-of-round code: the default line at N = 8 (C3 headline + C5 record), then C4 at N = 8
mkdir -p gpurun_out; nvidia-smi -L | wc -l
run() { n=$1; name=$2; shift 2; (NCCL_DEBUG=INFO NCCL_DEBUG_FILE=gpurun_out/${name}_nccl_%h_%p.log timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $n "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"); grep -h "NCCL INFO comm .* nranks\|Init COMPLETE" gpurun_out/${name}_nccl_* 2>/dev/null | head -2 | cut -c1-160; rm -f gpurun_out/${name}_nccl_*; }
run 8 fin_n8_bench --steps 20 --warmup 3
run 8 fin_n8_bench_c4 --steps 20 --warmup 3 --workload c4
python - <<'PY'
import json
for n in ("fin_n8_bench","fin_n8_bench_c4"):
    try:
        d=json.load(open("gpurun_out/%s.json"%n))
        print(n, "headline %.4f (%s) value %.1f seq %.4f repeats %.4f frames %.4f nccl %.4f e2e fps %.0f | %s"%(d["ms_per_step"], d["config"]["submission_choice"], d["value"], d["sequence_api"]["ms_per_step"], d["repeats_ms_per_step"]["median"], d["per_frame_api"]["ms_per_step"], d["nccl_gather_baseline"]["ms_per_step"], d["e2e"]["fps"], d["roofline"]["kernel"]), {k:v for k,v in d["verify"].items() if k!="what"}, "c5", (d.get("c5") or {}).get("ms_per_step"), (d.get("c5") or {}).get("kernel"), ((d.get("c5") or {}).get("verify") or {}).get("ok"))
    except Exception as e: print(n, "ERR", e)
PY
