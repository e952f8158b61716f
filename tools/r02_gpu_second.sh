#!/bin/bash
# round 2, second GPU pass: whole GPU suite, then bench lines for C3 and C4 with and without predicted leaps
mkdir -p gpurun_out; nvidia-smi -L | head -2
(timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r02b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_pytest.log); tail -14 gpurun_out/r02b_pytest.log
(timeout 500 python bench.py --steps 20 --warmup 3 > gpurun_out/r02b_bench_c3.json 2> gpurun_out/r02b_bench_c3.err; echo "bench rc=$?"); tail -3 gpurun_out/r02b_bench_c3.err
(timeout 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02b_bench_c4.json 2> gpurun_out/r02b_bench_c4.err; echo "c4 rc=$?"); tail -3 gpurun_out/r02b_bench_c4.err
(VR5_ESL_EXACT=1 timeout 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02b_bench_c4_exact.json 2> gpurun_out/r02b_bench_c4_exact.err; echo "c4 exact rc=$?")
python - <<'PY'
import json
for n in ("c3","c4","c4_exact"):
    try:
        d=json.load(open("gpurun_out/r02b_bench_%s.json"%n))
        print(n, "seq ms/frame %.4f"%d["ms_per_step"], "single %.4f"%d["roofline"]["single_frame_launch_ms"], "per-frame-api %.4f"%d["per_frame_api"]["ms_per_step"], "value %.1f"%d["value"], "e2e fps %.0f"%d["e2e"]["fps"], "verify", d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb"))
        if "policies" in d: print("  trif", d["policies"]["trilinear_fma"]["ms_per_step"], d["policies"]["trilinear_fma"]["frac"], d["policies"]["trilinear_fma"]["single_frame_launch_ms"])
        if "c5" in d: print("  c5", d["c5"]["ms_per_step"], d["c5"]["gsamples"], d["c5"]["verify"])
    except Exception as e: print(n, "ERR", e)
PY
