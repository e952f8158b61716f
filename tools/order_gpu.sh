#!/bin/sh
# tools/order_gpu.sh : on the GPU box - the whole GPU suite, then C3 bench lines with the centre-out and the row-major queue order
mkdir -p gpurun_out
t0=$(date +%s); timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_trilinear.py tests/test_peer_gather.py -m gpu -x -q > gpurun_out/order_pytest.log 2>&1
echo "pytest rc=$? in $(( $(date +%s) - t0 )) s"; tail -3 gpurun_out/order_pytest.log
for s in hybridw trilinear-fma; do
  for o in rows cols rings; do
    VR5_QUEUE_ORDER=$o python bench.py --steps 120 --warmup 5 --no-cpu-baseline --sampler $s > gpurun_out/order_${s}_$o.json 2> gpurun_out/order_${s}_$o.err
  done
done
for o in rows cols; do
  VR5_QUEUE_ORDER=$o python bench.py --workload c4 --steps 24 --warmup 3 --no-cpu-baseline > gpurun_out/order_c4_$o.json 2> gpurun_out/order_c4_$o.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/order_*.json")):
    try:
        d = json.load(open(f))
        print("#", f, "step %.4f ms  kernel %.4f ms  flushed %.4f  %.1f Gs/s  frac %.3f  e2e %.1f fps" % (d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_l2_flushed"], d["value"], d["roofline"]["frac"], d["e2e"]["fps"]))
    except Exception as e:
        print("#", f, "FAILED", e)
PY
