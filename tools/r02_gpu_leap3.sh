#!/bin/bash
# leaps over transparent bricks, third pass (brick-to-brick chains): parity, C4 per brick edge and without leaps
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_leap.py tests/test_zero_skip.py -m gpu -x -q 2>&1 | tail -15) > gpurun_out/leap3_pytest.log; tail -4 gpurun_out/leap3_pytest.log
B="--steps 20 --warmup 3 --no-cpu-baseline --no-policies --no-c5"
run() { name=$1; shift; envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  (timeout 400 env "${envs[@]}" python bench.py $B "$@" > gpurun_out/leap3_$name.json 2> gpurun_out/leap3_$name.err; echo "$name rc=$?"); }
run c4_default A=1 -- --workload c4
run c4_noleap VR5_NO_LEAP=1 -- --workload c4
run c4_b16 VR5_LEAP_BRICK=4 -- --workload c4
run c4_b64 VR5_LEAP_BRICK=6 -- --workload c4
run c4_b8 VR5_LEAP_BRICK=3 -- --workload c4
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/leap3_*.json")):
    try:
        d = json.load(open(f)); z = d.get("zero_skip") or {}
        print(f.split("/")[-1], "seq %.4f single %.4f per-frame-api %.4f | %s | verify %s/%s | leaps/frame %s left out %.3f" % (
            d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["roofline"]["kernel"],
            d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb"), z.get("leaps_per_frame"), z.get("fraction_of_samples_left_out", 0)))
    except Exception as e:
        print(f, "ERR", e)
PY
