#!/bin/bash
# zero-skip in the trilinear policies: parity (zero-skip + trilinear files, then the whole GPU suite), the reference's benchmark sweep
# through the contracted trilinear policy with the skip on / off, and the C3 policy line (semi table: plain kernel, must not move)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_zero_skip.py tests/test_trilinear.py tests/test_trilinear_pin.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/trizs_pytest.log; tail -3 gpurun_out/trizs_pytest.log
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8; echo "full rc=${PIPESTATUS[0]}") > gpurun_out/trizs_pytest_full.log; tail -3 gpurun_out/trizs_pytest_full.log
for cfg in default nozskip default2 nozskip2; do
  case $cfg in default*) E="A=1";; nozskip*) E="VR5_NO_ZSKIP=1";; esac
  (env $E timeout 400 python tools/volr_suite.py --sampler trilinear-fma --n 512 --window 1920 1080 --reps 5 --json gpurun_out/trizs_suite_$cfg.json > gpurun_out/trizs_suite_$cfg.log 2>&1; echo "suite $cfg rc=$?")
done
(timeout 500 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-c5 > gpurun_out/trizs_c3.json 2> gpurun_out/trizs_c3.err; echo "c3 rc=$?")
python - <<'PY'
import json, statistics
d=json.load(open("gpurun_out/trizs_c3.json")); p=d["policies"]["trilinear_fma"]
print("c3 seq %.4f | trif %.4f frac %.4f single %.4f %s"%(d["ms_per_step"], p["ms_per_step"], p["frac"], p["single_frame_launch_ms"], p["kernel"]))
rows={c:{r["config"]:r for r in json.load(open("gpurun_out/trizs_suite_%s.json"%c))["rows"]} for c in ("default","nozskip","default2","nozskip2")}
for k in rows["default"]:
    print("%-20s %-12s zero-skip %7.3f %7.3f | off %7.3f %7.3f" % (k, rows["default"][k]["volume"], rows["default"][k]["avg_ms"], rows["default2"][k]["avg_ms"], rows["nozskip"][k]["avg_ms"], rows["nozskip2"][k]["avg_ms"]))
g=lambda a: statistics.geometric_mean([min(rows[a][k]["avg_ms"], rows[a+"2"][k]["avg_ms"]) for k in rows["default"]])
print("geometric mean: zero-skip %.4f off %.4f"%(g("default"), g("nozskip")))
PY
