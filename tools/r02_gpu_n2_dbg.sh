#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/n2dbg_*
timeout 400 python -m pytest tests/test_sequence.py tests/test_peer_gather.py tests/test_hostring.py tests/test_volume_broadcast.py tests/test_bands.py tests/test_zz_no_alloc.py -m gpu -q -rs > gpurun_out/r02_n2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_n2_pytest.log; tail -4 gpurun_out/r02_n2_pytest.log
python tools/dbg_ring.py 2>&1 | grep -c ok
run() { name=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/$name.json 2> gpurun_out/$name.err || echo "$name FAILED"; }
run r02_n2_bench A=1
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_n2_bench.json"))
print("headline %.4f (%s) seq %.4f repeats-median %.4f per-frame-api %.4f nccl %.4f e2e fps %.0f"%(d["ms_per_step"], d["config"]["submission_choice"], d["sequence_api"]["ms_per_step"], d["repeats_ms_per_step"]["median"], d["per_frame_api"]["ms_per_step"], d["nccl_gather_baseline"]["ms_per_step"], d["e2e"]["fps"]))
print({k:v for k,v in d["verify"].items() if k!="what"}, d["c5"]["ms_per_step"], d["c5"]["verify"]["ok"])
PY
