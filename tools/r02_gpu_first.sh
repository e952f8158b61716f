#!/bin/bash
# round 2, first GPU pass: sequence tests + headless C++ driver, GPURenderer4 pin measurement, one bench line
mkdir -p gpurun_out; nvidia-smi -L | head -2
(timeout 300 python -m pytest tests/test_sequence.py tests/test_headless.py -m gpu -q -x > gpurun_out/r02a_seq.log 2>&1; echo "seq rc=$?" >> gpurun_out/r02a_seq.log); tail -15 gpurun_out/r02a_seq.log
(timeout 400 python tools/ref4_pin.py --full > gpurun_out/r02a_ref4_pin.json 2> gpurun_out/r02a_ref4_pin.err; echo "pin rc=$?"); tail -3 gpurun_out/r02a_ref4_pin.err
(timeout 500 python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"); tail -5 gpurun_out/r02a_bench.err; head -c 3000 gpurun_out/r02a_bench.json
