#!/bin/bash
# ncu: the per-rank kernel of an 8-GPU job, played by ONE GPU (ncu is single-process): rank's share as a centre band and as
# interleaved strips, 4-frame sequence launches
mkdir -p gpurun_out
for cfg in band-centre strips; do
  C="python tools/sim_rank.py --world 8 --steps 4 --only $cfg"
  $C > gpurun_out/ncu_sim_$cfg.json 2> gpurun_out/ncu_sim_$cfg.err && \
  ncu --set full --clock-control none -k regex:cast_kernel -s 5 -c 1 -f -o gpurun_out/r02_rank_of_8_$cfg $C > gpurun_out/ncu_sim_$cfg.log 2>&1; echo "$cfg ncu rc=$?"
done
ls -la gpurun_out/r02_rank_of_8_*.ncu-rep
