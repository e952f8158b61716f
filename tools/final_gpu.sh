#!/bin/sh
# tools/final_gpu.sh : end-of-round validation on one B200 - GPU suite, smoke, the default bench line, the reference arm, the
# other workloads / samplers, then the ncu launch list of the default bench command (after it exited 0 without ncu).
mkdir -p gpurun_out
t0=$(date +%s); timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1
echo "pytest rc=$? in $(( $(date +%s) - t0 )) s"; tail -2 gpurun_out/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/final_smoke.log
python bench.py > gpurun_out/final_bench_c3_n1.json 2> gpurun_out/final_bench_c3_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 6 --warmup 1 > gpurun_out/final_bench_c3_reference_arm.json 2> gpurun_out/final_ref.err; echo "reference arm rc=$?"
python bench.py --workload c4 --steps 24 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_c4_n1.json 2> gpurun_out/final_c4.err
python bench.py --workload c5 --steps 24 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_c5_n1.json 2> gpurun_out/final_c5.err
python bench.py --sampler trilinear --no-cpu-baseline > gpurun_out/final_bench_c3_trilinear_n1.json 2> gpurun_out/final_tri.err
python bench.py --sampler trilinear-fma --no-cpu-baseline > gpurun_out/final_bench_c3_trilinear_fma_n1.json 2> gpurun_out/final_trif.err
python bench.py --steps 60 --warmup 3 --no-cpu-baseline > gpurun_out/final_pre_ncu.json 2> gpurun_out/final_pre_ncu.err; echo "pre-ncu bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/final_launches_c3_default.csv python bench.py --steps 60 --warmup 3 --no-cpu-baseline > gpurun_out/final_ncu_list.log 2>&1; echo "ncu list rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/final_bench_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        r = d.get("roofline", {})
        print("#", f.split("final_bench_")[1], "step %.4f ms  %.1f Gs/s  fps %.1f  e2e fps %s  kernel %s  frac %s  in-stream %s" % (
            d["ms_per_step"], d["value"], d.get("fps", 0), d["e2e"].get("fps"), r.get("kernel_ms"), r.get("frac"), r.get("frac_in_stream")))
    except Exception as e:
        print("#", f, "FAILED", e)
PY
