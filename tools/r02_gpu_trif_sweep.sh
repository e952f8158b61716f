#!/bin/bash
mkdir -p gpurun_out
for cd in 6:2 6:1 5:2 4:2 4:3; do
  c=${cd%:*}; d=${cd#*:}
  VR5_TRIF_CTAS=$c VR5_TRIF_DEPTH=$d timeout 200 python bench.py --steps 40 --warmup 3 --no-cpu-baseline --no-c5 --no-policies --sampler trilinear-fma > gpurun_out/trifsweep_c${c}_d${d}.json 2> gpurun_out/trifsweep_c${c}_d${d}.err || echo "$cd FAILED"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/trifsweep_*.json")):
    try:
        d=json.load(open(f)); print(f.split("/")[-1], "seq %.4f frames %.4f single %.4f frac %.4f verify %s/%s"%(d["sequence_api"]["ms_per_step"], d["per_frame_api"]["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["roofline"]["frac"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb")))
    except Exception as e: print(f,"ERR",e)
PY
