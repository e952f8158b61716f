#!/bin/sh
# tools/sweep.sh OUT "variant:sampler:ENV=VALUE ..." (variant cur = the built libvr5.so) : on the GPU box, run a short C3 bench per combination and print one line each
out=$1; shift
for spec in $*; do
  v=$(echo $spec | cut -d: -f1); s=$(echo $spec | cut -d: -f2); g=$(echo $spec | cut -d: -f3)
  if [ "$v" != cur ]; then cp volume-rendering_b200/build/var/$v.so volume-rendering_b200/libvr5.so; fi
  env $g python bench.py --steps 120 --warmup 5 --no-cpu-baseline --sampler $s > gpurun_out/${out}_$v-$s-$g.json 2> gpurun_out/${out}_$v-$s-$g.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${out}_$v-$s-$g.json"))
    print("$spec", "step %.4f ms  kernel %.4f ms  flushed %.4f  %.1f Gs/s" % (d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_l2_flushed"], d["value"]))
except Exception as e:
    print("$spec FAILED", e)
PY
done
