#!/bin/bash
# two-GPU lease: the multi-process tests that skip on one GPU (0 skips expected here) and a 2-rank bench line
mkdir -p gpurun_out; nvidia-smi -L
(timeout 600 python -m pytest tests/test_peer_gather.py tests/test_hostring.py tests/test_sequence.py tests/test_volume_broadcast.py -m gpu -q -rs > gpurun_out/r02_n2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_n2_pytest.log); tail -8 gpurun_out/r02_n2_pytest.log
(NCCL_DEBUG=INFO NCCL_DEBUG_FILE=gpurun_out/r02_n2_nccl_%h_%p.log timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02_n2_bench.json 2> gpurun_out/r02_n2_bench.err; echo "bench rc=$?"); tail -5 gpurun_out/r02_n2_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_n2_bench.json"))
print("n2 seq ms/frame %.4f"%d["ms_per_step"], "value %.1f"%d["value"], "per-frame-api %.4f"%d["per_frame_api"]["ms_per_step"], "nccl %.4f"%d["nccl_gather_baseline"]["ms_per_step"], "e2e fps %.0f"%d["e2e"]["fps"])
print(" verify", d["verify"]); print(" kernel_vs_ideal", d["kernel_vs_ideal"]); print(" c5", d.get("c5"))
PY
