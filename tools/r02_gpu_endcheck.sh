#!/bin/bash
# end-of-round check with the driver's own commands, then the ncu launch list of the same bench command (VR5_RING_SLOTS=32: DESIGN section 6)
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -4; echo "pytest rc=${PIPESTATUS[0]}") > gpurun_out/end_pytest.log; tail -2 gpurun_out/end_pytest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1)
(timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/end_reference.json 2> gpurun_out/end_reference.err; echo "reference rc=$?")
(timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/end_bench.json 2> gpurun_out/end_bench.err; echo "bench rc=$?")
export VR5_RING_SLOTS=32
D="python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --c5-frames 2"
timeout 300 $D > gpurun_out/end_ncu_plain.json 2> gpurun_out/end_ncu_plain.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches_final.csv $D > gpurun_out/end_ncu.log 2>&1; echo "launch list rc=$?"
python - <<'PY'
import json
r=json.load(open("gpurun_out/end_reference.json")); d=json.load(open("gpurun_out/end_bench.json"))
print("reference", r["value"], r["unit"], r["fps"], "| bench value %.1f e2e %.1f fps %.0f ms %.4f frac %.4f launches %s"%(d["value"], d["e2e"]["value"], d["e2e"]["fps"], d["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"]))
print(" ratio e2e %.0f device %.0f"%(d["e2e"]["value"]/r["value"], d["value"]/r["value"]), "| trif frac", d["policies"]["trilinear_fma"]["frac"], "| c5", d["c5"]["ms_per_step"], d["c5"]["verify"]["ok"], "| verify", {k:v for k,v in d["verify"].items() if k!="what"}, d["clocks"])
PY
