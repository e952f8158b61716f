#!/bin/bash
# zero-skip kernels, second pass: parity (new file + whole GPU suite), C4 lines (skip on / off, same box), ncu captures of the C4
# zero-skip sequence launch and of the C5 wide-window launch (each after its command has exited 0 without ncu), final N = 1 lines
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_zero_skip.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/zs2_pytest_zs.log; tail -2 gpurun_out/zs2_pytest_zs.log
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8; echo "full rc=${PIPESTATUS[0]}") > gpurun_out/zs2_pytest_full.log; tail -3 gpurun_out/zs2_pytest_full.log
(timeout 500 python bench.py --workload c4 --steps 20 --warmup 3 --no-policies --no-c5 > gpurun_out/zs2_c4.json 2> gpurun_out/zs2_c4.err; echo "c4 rc=$?")
(VR5_NO_ZSKIP=1 timeout 500 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu-baseline --no-policies --no-c5 > gpurun_out/zs2_c4_nozs.json 2> gpurun_out/zs2_c4_nozs.err; echo "c4 nozs rc=$?")
(timeout 700 python bench.py --steps 20 --warmup 3 > gpurun_out/zs2_n1_bench.json 2> gpurun_out/zs2_n1_bench.err; echo "n1 rc=$?")
(timeout 700 python bench.py --steps 360 --warmup 8 --no-c5 > gpurun_out/zs2_n1_bench_360.json 2> gpurun_out/zs2_n1_bench_360.err; echo "n1 360 rc=$?")
B="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-policies --no-c5"
timeout 300 $B --workload c4 > gpurun_out/zs2_ncu_c4_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cast_kernel -s 5 -c 1 -f -o gpurun_out/r02_c4_zs_seq $B --workload c4 > gpurun_out/zs2_ncu_c4.log 2>&1; echo "c4 ncu rc=$?"
timeout 400 $B --workload c5 > gpurun_out/zs2_ncu_c5_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:cast_kernel -s 5 -c 1 -f -o gpurun_out/r02_c5_wide_seq $B --workload c5 > gpurun_out/zs2_ncu_c5.log 2>&1; echo "c5 ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/zs2_*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "seq %.4f single %.4f per-frame-api %.4f e2e %.1f blocking %s | %s | verify %s/%s | frac %.4f" % (
            d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["e2e"]["value"],
            d.get("e2e_blocking", {}).get("value"), d["roofline"]["kernel"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb"), d["roofline"]["frac"]))
        if "c5" in d: print("   c5", d["c5"]["ms_per_step"], d["c5"]["gsamples"], d["c5"]["kernel"], d["c5"]["verify"].get("ok"))
        if "policies" in d: print("   trif", d["policies"]["trilinear_fma"]["ms_per_step"], d["policies"]["trilinear_fma"]["frac"])
    except Exception as e:
        print(f, "ERR", e)
PY
