#!/bin/sh
# tools/trif_quick.sh : on the GPU box - trilinear parity tests, one C3 bench line of the default contracted march, one ncu capture
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_trilinear.py -m gpu -x -q > gpurun_out/trifq_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/trifq_pytest.log
python bench.py --steps 360 --warmup 8 --no-cpu-baseline --sampler trilinear-fma > gpurun_out/trifq_bench.json 2> gpurun_out/trifq_bench.err
python -c "
import json; d=json.load(open('gpurun_out/trifq_bench.json'))
print('step %.4f ms  kernel %.4f ms  flushed %.4f  %.1f Gs/s  frac %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['kernel_ms_l2_flushed'], d['value'], d['roofline']['frac']))"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:tri_kernel -s 30 -c 2 -f -o gpurun_out/trifq python bench.py --steps 12 --warmup 3 --no-cpu-baseline --sampler trilinear-fma > gpurun_out/trifq_ncu.log 2>&1
echo "ncu rc=$?"
