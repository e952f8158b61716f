#!/bin/bash
# ncu pass (one gpurun call): full captures of the sequence launch on C4, C3 (nearest) and C3 (trilinear-fma), then the launch list of
# the default bench command. Each command has exited 0 without ncu right before (&&).
mkdir -p gpurun_out
B="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-policies --no-c5"
$B --workload c4 > gpurun_out/ncu_c4_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:cast_kernel -s 5 -c 1 -f -o gpurun_out/r02_c4_seq $B --workload c4 > gpurun_out/ncu_c4.log 2>&1; echo "c4 ncu rc=$?"
$B --workload c3 > gpurun_out/ncu_c3_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:cast_kernel -s 5 -c 1 -f -o gpurun_out/r02_c3_seq $B --workload c3 > gpurun_out/ncu_c3.log 2>&1; echo "c3 ncu rc=$?"
$B --workload c3 --sampler trilinear-fma > gpurun_out/ncu_trif_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:tri_kernel -s 5 -c 1 -f -o gpurun_out/r02_c3_trif_seq $B --workload c3 --sampler trilinear-fma > gpurun_out/ncu_trif.log 2>&1; echo "trif ncu rc=$?"
D="python bench.py --steps 20 --warmup 3"
$D > gpurun_out/ncu_default_plain.json 2> gpurun_out/ncu_default_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv $D > gpurun_out/ncu_default.log 2>&1; echo "launch list rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_launches.csv
