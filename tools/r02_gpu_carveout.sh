#!/bin/bash
# shared-memory carve-out preference of the persistent kernels (more L1): C3 line (nearest + trilinear-fma policy + c5) and C4 with the
# preference and with the driver's default, same box
mkdir -p gpurun_out
for cfg in carve nocarve carve2 nocarve2; do
  case $cfg in carve*) E="A=1";; nocarve*) E="VR5_NO_CARVEOUT=1";; esac
  (env $E timeout 500 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --c5-frames 4 > gpurun_out/co_${cfg}_c3.json 2> gpurun_out/co_${cfg}_c3.err; echo "$cfg c3 rc=$?")
done
for cfg in carve nocarve; do
  case $cfg in carve*) E="A=1";; nocarve*) E="VR5_NO_CARVEOUT=1";; esac
  (env $E timeout 500 python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline --no-policies --no-c5 > gpurun_out/co_${cfg}_c4.json 2> gpurun_out/co_${cfg}_c4.err; echo "$cfg c4 rc=$?")
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/co_*.json")):
    try:
        d=json.load(open(f))
        print(f.split("/")[-1], "seq %.4f single %.4f frames %.4f e2e fps %.0f blocking %s verify %s/%s"%(d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["e2e"]["fps"], d.get("e2e_blocking",{}).get("fps"), d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb")),
              "| c5", (d.get("c5") or {}).get("ms_per_step"), "| trif", (d.get("policies") or {}).get("trilinear_fma",{}).get("ms_per_step"), (d.get("policies") or {}).get("trilinear_fma",{}).get("frac"), (d.get("policies") or {}).get("trilinear_fma",{}).get("single_frame_launch_ms"))
    except Exception as e: print(f, "ERR", e)
PY
