#!/bin/sh
# tools/pf_gpu.sh : on the GPU box - parity suite with the built library, then C3 / C4 / C2 bench lines against the pf1 variant build
mkdir -p gpurun_out
t0=$(date +%s); timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pf_pytest.log 2>&1
echo "pytest rc=$? in $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/pf_pytest.log
cp volume-rendering_b200/libvr5.so /tmp/cur.so
for v in cur pf1 cur pf1; do
  if [ $v = cur ]; then cp /tmp/cur.so volume-rendering_b200/libvr5.so; else cp volume-rendering_b200/build/var/$v.so volume-rendering_b200/libvr5.so; fi
  n=$(ls gpurun_out/pf_c3_${v}_*.json 2>/dev/null | wc -l)
  python bench.py --steps 240 --warmup 5 --no-cpu-baseline > gpurun_out/pf_c3_${v}_$n.json 2> gpurun_out/pf_c3_${v}_$n.err
  python bench.py --workload c4 --steps 24 --warmup 3 --no-cpu-baseline > gpurun_out/pf_c4_${v}_$n.json 2> gpurun_out/pf_c4_${v}_$n.err
done
cp /tmp/cur.so volume-rendering_b200/libvr5.so
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/pf_c*.json")):
    try:
        d = json.load(open(f))
        print("#", f, "step %.4f ms  kernel %.4f ms  flushed %.4f  %.1f Gs/s  e2e %.1f fps" % (d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_l2_flushed"], d["value"], d["e2e"]["fps"]))
    except Exception as e:
        print("#", f, "FAILED", e)
PY
