#!/bin/bash
# C5 (2048^3, 7680x4320): tile-queue orders (never measured at this size: DRAM traffic is 2x the compulsory footprint, L2 hit 46 %)
mkdir -p gpurun_out
for o in rows cols rings; do
  (VR5_QUEUE_ORDER=$o timeout 300 python bench.py --workload c5 --steps 8 --warmup 3 --no-cpu-baseline --no-policies --no-c5 > gpurun_out/c5o_$o.json 2> gpurun_out/c5o_$o.err; echo "$o rc=$?")
done
python - <<'PY'
import json
for o in ("rows","cols","rings"):
    try:
        d=json.load(open("gpurun_out/c5o_%s.json"%o))
        print(o, "seq %.3f single %.3f frames %.3f verify %s/%s %s"%(d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb"), d["roofline"]["kernel"]))
    except Exception as e: print(o, "ERR", e)
PY
