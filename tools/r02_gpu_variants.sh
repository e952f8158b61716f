#!/bin/bash
# experiment sweep: C3 / C4 bench lines per library variant (volume-rendering_b200/build/var/*.so copied over libvr5.so)
mkdir -p gpurun_out; cp volume-rendering_b200/libvr5.so /tmp/libvr5_default.so
for v in default "$@"; do
  if [ "$v" = default ]; then cp /tmp/libvr5_default.so volume-rendering_b200/libvr5.so; else cp volume-rendering_b200/build/var/$v.so volume-rendering_b200/libvr5.so; fi
  for wl in c3 c4; do
    timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline --no-policies --no-c5 > gpurun_out/var_${v}_${wl}.json 2> gpurun_out/var_${v}_${wl}.err || echo "$v $wl FAILED"
  done
done
cp /tmp/libvr5_default.so volume-rendering_b200/libvr5.so
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/var_*.json")):
    try:
        d=json.load(open(f)); print(f.split("/")[-1], "seq %.4f single %.4f per-frame-api %.4f verify %s/%s"%(d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb")))
    except Exception as e: print(f, "ERR", e)
PY
