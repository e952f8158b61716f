#!/bin/bash
# ncu pass 2 (one gpurun call): the trilinear-fma sequence launch (full set) and the launch list of the default bench command.
# VR5_RING_SLOTS=32: ncu serialises kernels, so a sequence kernel must never have to wait for its consumer.
mkdir -p gpurun_out
export VR5_RING_SLOTS=32
B="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-policies --no-c5 --sampler trilinear-fma"
$B > gpurun_out/ncu_trif_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:tri_kernel -s 5 -c 1 -f -o gpurun_out/r02_c3_trif_seq $B > gpurun_out/ncu_trif.log 2>&1; echo "trif ncu rc=$?"
D="python bench.py --steps 20 --warmup 3 --no-c5"
$D > gpurun_out/ncu_default_plain.json 2> gpurun_out/ncu_default_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_launches.csv $D > gpurun_out/ncu_default.log 2>&1; echo "launch list rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_launches.csv; tail -3 gpurun_out/ncu_default.log
