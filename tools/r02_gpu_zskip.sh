#!/bin/bash
# zero-skip kernels (ZS): parity (new test file, then the whole GPU suite), C4 with the skip on / off in the same run, prefetch depth 2
# under the skip (build/var/pf2.so), the default-TF C2 case, then the default N = 1 line with the final code
mkdir -p gpurun_out; cp volume-rendering_b200/libvr5.so /tmp/libvr5_default.so
(timeout 900 python -m pytest tests/test_zero_skip.py -m gpu -x -q 2>&1 | tail -5) > gpurun_out/zs_pytest_zs.log; tail -3 gpurun_out/zs_pytest_zs.log
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8; echo "full rc=${PIPESTATUS[0]}") > gpurun_out/zs_pytest_full.log; tail -3 gpurun_out/zs_pytest_full.log
B="--steps 20 --warmup 3 --no-cpu-baseline --no-policies --no-c5"
# run NAME ENV... -- bench args
run() { name=$1; shift; envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  (timeout 400 env "${envs[@]}" python bench.py $B "$@" > gpurun_out/zs_$name.json 2> gpurun_out/zs_$name.err; echo "$name rc=$?"); }
run c4_zs        A=1 -- --workload c4
run c4_nozs      VR5_NO_ZSKIP=1 -- --workload c4
run c4_zs_window A=1 -- --workload c4 --sampler window
cp volume-rendering_b200/build/var/pf2.so volume-rendering_b200/libvr5.so
run c4_zs_pf2        A=1 -- --workload c4
run c4_zs_pf2_window A=1 -- --workload c4 --sampler window
run c3_pf2_check     A=1 -- --workload c3
cp /tmp/libvr5_default.so volume-rendering_b200/libvr5.so
(timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/zs_n1_bench.json 2> gpurun_out/zs_n1_bench.err; echo "n1 rc=$?")
(timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/zs_n1_reference.json 2> gpurun_out/zs_n1_reference.err; echo "ref rc=$?")
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/zs_smoke.log 2>&1; echo "smoke rc=$?"); tail -2 gpurun_out/zs_smoke.log
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/zs_*.json")):
    try:
        d = json.load(open(f))
        if d.get("impl") == "reference":
            print(f.split("/")[-1], "reference", d["value"], d["unit"]); continue
        print(f.split("/")[-1], "seq %.4f single %.4f per-frame-api %.4f e2e %.1f blocking %s | %s | verify %s/%s" % (
            d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["e2e"]["value"],
            d.get("e2e_blocking", {}).get("value"), d["roofline"]["kernel"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb")))
        if "c5" in d: print("   c5", d["c5"]["ms_per_step"], d["c5"]["gsamples"], d["c5"]["kernel"], d["c5"]["verify"].get("ok"))
        if "policies" in d: print("   trif", d["policies"]["trilinear_fma"]["ms_per_step"], d["policies"]["trilinear_fma"]["frac"])
    except Exception as e:
        print(f, "ERR", e)
PY
