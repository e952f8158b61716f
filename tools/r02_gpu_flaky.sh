#!/bin/bash
mkdir -p gpurun_out
echo "== forced per-CTA reads (default)"; python tools/dbg_ring.py 2>&1 | grep -c " ok"
echo "== per-warp reads allowed under the hand-shake (VR5_SEQ_DEBUG=8)"
VR5_SEQ_DEBUG=8 python tools/dbg_ring.py 2>&1 | tail -10
VR5_SEQ_DEBUG=8 VR5_WARP_FETCH=1 python tools/dbg_ring.py 2>&1 | tail -10
VR5_SEQ_DEBUG=8 timeout 300 python -m pytest tests/test_sequence.py -m gpu -q 2>&1 | tail -2
timeout 300 python -m pytest tests/test_sequence.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2
for v in A=1 VR5_SEQ_DEBUG=8; do env $v timeout 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-c5 --no-policies --no-cpu-baseline --submission sequence 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); print('$v c4 seq %.4f frames %.4f verify %s/%s'%(d['sequence_api']['ms_per_step'], d['per_frame_api']['ms_per_step'], d['verify']['sequence_frames_equal'], d['verify']['max_lsb']))"; done
