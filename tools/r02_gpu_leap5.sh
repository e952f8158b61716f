#!/bin/bash
# leaps gated on eight transparent samples in a row, brick edge >= 32: parity files, C4, and the reference's benchmark sweep with leaps on / off
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_leap.py tests/test_zero_skip.py -m gpu -x -q 2>&1 | tail -15) > gpurun_out/leap5_pytest.log; tail -3 gpurun_out/leap5_pytest.log
B="--steps 20 --warmup 3 --no-cpu-baseline --no-policies --no-c5"
(timeout 400 python bench.py $B --workload c4 > gpurun_out/leap5_c4.json 2> gpurun_out/leap5_c4.err; echo "c4 rc=$?")
(VR5_NO_LEAP=1 timeout 400 python bench.py $B --workload c4 > gpurun_out/leap5_c4_noleap.json 2> gpurun_out/leap5_c4_noleap.err; echo "c4 noleap rc=$?")
for cfg in default noleap nozskip default2 noleap2; do
  case $cfg in default*) E="A=1";; noleap*) E="VR5_NO_LEAP=1";; nozskip) E="VR5_NO_ZSKIP=1";; esac
  (env $E timeout 400 python tools/volr_suite.py --n 512 --window 1920 1080 --reps 5 --json gpurun_out/leap5_suite_$cfg.json > gpurun_out/leap5_suite_$cfg.log 2>&1; echo "suite $cfg rc=$?")
done
python - <<'PY'
import json
for n in ("leap5_c4","leap5_c4_noleap"):
    d=json.load(open("gpurun_out/%s.json"%n)); z=d.get("zero_skip") or {}
    print(n, "seq %.4f single %.4f frames %.4f verify %s/%s left out %.3f"%(d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb"), z.get("fraction_of_samples_left_out",0)))
rows={}
for cfg in ("default","noleap","nozskip","default2","noleap2"):
    rows[cfg]={r["config"]:r for r in json.load(open("gpurun_out/leap5_suite_%s.json"%cfg))["rows"]}
for k in rows["default"]:
    print("%-20s %-12s default %7.3f %7.3f | no leaps %7.3f %7.3f | no zero-skip %7.3f" % (k, rows["default"][k]["volume"], rows["default"][k]["avg_ms"], rows["default2"][k]["avg_ms"], rows["noleap"][k]["avg_ms"], rows["noleap2"][k]["avg_ms"], rows["nozskip"][k]["avg_ms"]))
PY
