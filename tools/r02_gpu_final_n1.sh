#!/bin/bash
# end-of-round N = 1 lines (C3 default incl. c5 + policies + cpu baseline, 360 steps, C4, reference arm) and the reference's own benchmark
# sweep (default transfer function) with the zero-skip kernels / leaps on and off
mkdir -p gpurun_out
(timeout 700 python bench.py --steps 20 --warmup 3 > gpurun_out/fin_n1_bench.json 2> gpurun_out/fin_n1_bench.err; echo "n1 rc=$?")
(timeout 700 python bench.py --steps 360 --warmup 8 --no-c5 > gpurun_out/fin_n1_bench_360.json 2> gpurun_out/fin_n1_bench_360.err; echo "n1 360 rc=$?")
(timeout 500 python bench.py --workload c4 --steps 20 --warmup 3 --no-policies --no-c5 > gpurun_out/fin_n1_bench_c4.json 2> gpurun_out/fin_n1_bench_c4.err; echo "c4 rc=$?")
(timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/fin_n1_reference.json 2> gpurun_out/fin_n1_reference.err; echo "ref rc=$?")
for cfg in default noleap nozskip; do
  case $cfg in default) E="A=1";; noleap) E="VR5_NO_LEAP=1";; nozskip) E="VR5_NO_ZSKIP=1";; esac
  (env $E timeout 400 python tools/volr_suite.py --n 512 --window 1920 1080 --reps 5 --json gpurun_out/fin_suite_$cfg.json > gpurun_out/fin_suite_$cfg.log 2>&1; echo "suite $cfg rc=$?")
done
python - <<'PY'
import json
for n in ("fin_n1_bench","fin_n1_bench_360","fin_n1_bench_c4"):
    try:
        d=json.load(open("gpurun_out/%s.json"%n))
        print(n, "seq %.4f single %.4f frames %.4f e2e %.1f fps %.0f blocking %s | %s | verify %s/%s | frac %.4f"%(d["ms_per_step"], d["roofline"]["single_frame_launch_ms"], d["per_frame_api"]["ms_per_step"], d["e2e"]["value"], d["e2e"]["fps"], d.get("e2e_blocking",{}).get("fps"), d["roofline"]["kernel"], d["verify"].get("sequence_frames_equal"), d["verify"].get("max_lsb"), d["roofline"]["frac"]))
        if "c5" in d: print("   c5", d["c5"]["ms_per_step"], d["c5"]["gsamples"], d["c5"]["kernel"], d["c5"]["verify"].get("ok"))
        if "policies" in d: print("   trif", d["policies"]["trilinear_fma"]["ms_per_step"], d["policies"]["trilinear_fma"]["frac"])
        if "cpu_baseline" in d: print("   cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
    except Exception as e: print(n, "ERR", e)
try:
    print(json.load(open("gpurun_out/fin_n1_reference.json")))
except Exception as e: print("ref ERR", e)
rows={}
for cfg in ("default","noleap","nozskip"):
    try: rows[cfg]={r["config"]:r for r in json.load(open("gpurun_out/fin_suite_%s.json"%cfg))["rows"]}
    except Exception as e: print(cfg,"ERR",e)
if len(rows)==3:
    for k in rows["default"]:
        print("%-20s %-10s default %7.3f ms  no leaps %7.3f  no zero-skip %7.3f" % (k, rows["default"][k]["volume"], rows["default"][k]["avg_ms"], rows["noleap"][k]["avg_ms"], rows["nozskip"][k]["avg_ms"]))
PY
