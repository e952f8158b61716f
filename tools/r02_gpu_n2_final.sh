#!/bin/bash
# two-GPU lease with the end-of-round code: the multi-process tests (0 skips expected), the N = 2 line (C3 + C5 record), C4 at N = 2
mkdir -p gpurun_out; nvidia-smi -L
(timeout 600 python -m pytest tests/test_peer_gather.py tests/test_hostring.py tests/test_sequence.py tests/test_volume_broadcast.py -m gpu -q -rs > gpurun_out/fin_n2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/fin_n2_pytest.log); tail -4 gpurun_out/fin_n2_pytest.log
run() { name=$1; shift; (timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"); tail -2 gpurun_out/$name.err; }
run fin_n2_bench --steps 20 --warmup 3
run fin_n2_bench_c4 --steps 20 --warmup 3 --workload c4
python - <<'PY'
import json
for n in ("fin_n2_bench", "fin_n2_bench_c4"):
    try:
        d=json.load(open("gpurun_out/%s.json"%n))
        print(n, "headline %.4f (%s) value %.1f seq %.4f frames %.4f nccl %.4f e2e fps %.0f | %s"%(d["ms_per_step"], d["config"]["submission_choice"], d["value"], d["sequence_api"]["ms_per_step"], d["per_frame_api"]["ms_per_step"], d["nccl_gather_baseline"]["ms_per_step"], d["e2e"]["fps"], d["roofline"]["kernel"]))
        print("  verify", {k:v for k,v in d["verify"].items() if k!="what"}, "c5", (d.get("c5") or {}).get("ms_per_step"), (d.get("c5") or {}).get("kernel"), ((d.get("c5") or {}).get("verify") or {}).get("ok"), "zs", {k:v for k,v in (d.get("zero_skip") or {}).items() if k!="note"})
    except Exception as e: print(n, "ERR", e)
PY
