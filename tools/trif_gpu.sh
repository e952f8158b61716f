#!/bin/sh
# tools/trif_gpu.sh : on the GPU box - parity tests of both trilinear policies, a C3 bench line per register-budget variant of
# the contracted march, then ONE ncu --set full capture of the fastest (after its bench command exited 0 without ncu).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_trilinear.py -m gpu -x -q > gpurun_out/trif_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/trif_pytest.log
for cd in 6:1 6:2 5:2 4:2 4:3 4:4; do
  c=${cd%:*}; d=${cd#*:}
  VR5_TRIF_CTAS=$c VR5_TRIF_DEPTH=$d python bench.py --steps 120 --warmup 5 --no-cpu-baseline --sampler trilinear-fma > gpurun_out/trif_c${c}_d${d}.json 2> gpurun_out/trif_c${c}_d${d}.err
done
VR5_WARP_FETCH=1 VR5_TRIF_CTAS=6 VR5_TRIF_DEPTH=2 python bench.py --steps 120 --warmup 5 --no-cpu-baseline --sampler trilinear-fma > gpurun_out/trif_c6_d2_warpfetch.json 2> gpurun_out/trif_c6_d2_warpfetch.err
python bench.py --steps 120 --warmup 5 --no-cpu-baseline --sampler trilinear > gpurun_out/trif_old_trilinear.json 2> gpurun_out/trif_old_trilinear.err
best=$(python - <<'PY'
import json, glob
b = None
for f in sorted(glob.glob("gpurun_out/trif_*.json")):
    try:
        d = json.load(open(f))
        print("#", f, "step %.4f ms  kernel %.4f ms  flushed %.4f  %.1f Gs/s  frac %.3f" % (d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_l2_flushed"], d["value"], d["roofline"]["frac"]), file=__import__("sys").stderr)
        if "_c" in f and "warpfetch" not in f and (b is None or d["ms_per_step"] < b[0]):
            c, dd = f.split("trif_c")[1].split(".json")[0].split("_d")
            b = (d["ms_per_step"], c, dd)
    except Exception as e:
        print("#", f, "FAILED", e, file=__import__("sys").stderr)
print("%s %s" % (b[1], b[2]) if b else "6 1")
PY
)
bc=${best% *}; bd=${best#* }
echo "best VR5_TRIF_CTAS=$bc VR5_TRIF_DEPTH=$bd"
VR5_TRIF_CTAS=$bc VR5_TRIF_DEPTH=$bd timeout 400 ncu --set full --clock-control none --import-source on -k regex:tri_kernel -s 30 -c 2 -f -o gpurun_out/trif_best \
  python bench.py --steps 12 --warmup 3 --no-cpu-baseline --sampler trilinear-fma > gpurun_out/trif_ncu.log 2>&1
echo "ncu rc=$? (VR5_TRIF_CTAS=$bc VR5_TRIF_DEPTH=$bd)"; echo "$bc $bd" > gpurun_out/trif_best_ctas.txt
