#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_trilinear.py tests/test_trilinear_pin.py tests/test_sequence.py -m gpu -q 2>&1 | tail -2
for i in 1 2; do timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-c5 2>/dev/null | python -c "import json,sys; d=json.load(sys.stdin); p=d['policies']['trilinear_fma']; print('c3 %.4f | trif %.4f frac %.4f single %.4f'%(d['ms_per_step'], p['ms_per_step'], p['frac'], p['single_frame_launch_ms']))"; done
