#!/bin/bash
# height of the tile-queue groups (32 / 64 / 128 pixels): C5 and C3 per library variant (build/var/gy4.so, gy5.so)
mkdir -p gpurun_out; cp volume-rendering_b200/libvr5.so /tmp/libvr5_default.so
for v in default gy4 gy5; do
  if [ "$v" = default ]; then cp /tmp/libvr5_default.so volume-rendering_b200/libvr5.so; else cp volume-rendering_b200/build/var/$v.so volume-rendering_b200/libvr5.so; fi
  (timeout 200 python bench.py --workload c5 --steps 6 --warmup 3 --no-cpu-baseline --no-policies --no-c5 > gpurun_out/grp_${v}_c5.json 2> gpurun_out/grp_${v}_c5.err; echo "$v c5 rc=$?")
  (timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c5 > gpurun_out/grp_${v}_c3.json 2> gpurun_out/grp_${v}_c3.err; echo "$v c3 rc=$?")
done
cp /tmp/libvr5_default.so volume-rendering_b200/libvr5.so
python - <<'PY'
import json
for v in ("default","gy4","gy5"):
    for w in ("c5","c3"):
        try:
            d=json.load(open("gpurun_out/grp_%s_%s.json"%(v,w)))
            print(v, w, "seq %.4f single %.4f frames %.4f verify %s/%s"%(d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb")), "trif", (d.get("policies") or {}).get("trilinear_fma",{}).get("ms_per_step"))
        except Exception as e: print(v, w, "ERR", e)
PY
