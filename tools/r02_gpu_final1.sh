#!/bin/bash
# one GPU: whole GPU suite, the default bench line, the reference arm, C4 line
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_full.log 2>&1; echo "full rc=$?" >> gpurun_out/r02_pytest_full.log); tail -3 gpurun_out/r02_pytest_full.log
(timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r02_n1_bench.json 2> gpurun_out/r02_n1_bench.err; echo "bench rc=$?"); tail -2 gpurun_out/r02_n1_bench.err
(timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_n1_reference.json 2> gpurun_out/r02_n1_reference.err; echo "ref rc=$?")
(timeout 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu-baseline --no-c5 > gpurun_out/r02_n1_bench_c4.json 2> gpurun_out/r02_n1_bench_c4.err; echo "c4 rc=$?")
(timeout 300 python bench.py --steps 360 --warmup 8 --no-cpu-baseline --no-c5 > gpurun_out/r02_n1_bench_360.json 2> gpurun_out/r02_n1_bench_360.err; echo "360 rc=$?")
python - <<'PY'
import json
for n in ("r02_n1_bench","r02_n1_bench_c4","r02_n1_bench_360"):
    try:
        d=json.load(open("gpurun_out/%s.json"%n))
        print(n, "headline %.4f (%s) seq %.4f frames %.4f single %.4f value %.1f frac %.4f e2e fps %.0f blocking fps %.0f verify %s/%s"%(d["ms_per_step"], d["config"]["submission_choice"], d["sequence_api"]["ms_per_step"], d["per_frame_api"]["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["value"], d["roofline"]["frac"], d["e2e"]["fps"], d["e2e_blocking"]["fps"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb")))
        if "policies" in d: p=d["policies"]["trilinear_fma"]; print("  trif ms %.4f frac %.4f single %.4f single-frac %.4f"%(p["ms_per_step"], p["frac"], p["single_frame_launch_ms"], p["single_frame_launch_frac"]))
        if "c5" in d: print("  c5", d["c5"]["ms_per_step"], d["c5"]["gsamples"], d["c5"]["verify"])
        if "cpu_baseline" in d: print("  cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
    except Exception as e: print(n, "ERR", e)
try: print(open("gpurun_out/r02_n1_reference.json").read()[:400])
except Exception as e: print(e)
PY
