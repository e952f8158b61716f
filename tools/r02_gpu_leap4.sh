#!/bin/bash
# leaps: whole GPU suite, then an ncu capture (full set, source view) of the C4 sequence launch with leaps
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8; echo "full rc=${PIPESTATUS[0]}") > gpurun_out/leap4_pytest_full.log; tail -3 gpurun_out/leap4_pytest_full.log
B="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-policies --no-c5"
timeout 300 $B --workload c4 > gpurun_out/leap4_ncu_c4_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cast_kernel -s 5 -c 1 -f -o gpurun_out/r02_c4_leap_seq $B --workload c4 > gpurun_out/leap4_ncu_c4.log 2>&1; echo "c4 ncu rc=$?"
ls -la gpurun_out/r02_c4_leap_seq.ncu-rep
