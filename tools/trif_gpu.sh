#!/bin/sh
# tools/trif_gpu.sh : on the GPU box - parity tests of both trilinear policies, a C3 bench line per register-budget variant of
# the contracted march, then ONE ncu --set full capture of the fastest (after its bench command exited 0 without ncu).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_trilinear.py -m gpu -x -q > gpurun_out/trif_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/trif_pytest.log
sh tools/sweep.sh trif "cur:trilinear:X=0 cur:trilinear-fma:VR5_TRIF_CTAS=6 cur:trilinear-fma:VR5_TRIF_CTAS=5 cur:trilinear-fma:VR5_TRIF_CTAS=4"
best=$(python - <<'PY'
import json
b = None
for c in (6, 5, 4):
    try:
        d = json.load(open("gpurun_out/trif_cur-trilinear-fma-VR5_TRIF_CTAS=%d.json" % c))
        if b is None or d["ms_per_step"] < b[0]:
            b = (d["ms_per_step"], c)
    except Exception:
        pass
print(b[1] if b else 6)
PY
)
echo "best VR5_TRIF_CTAS=$best"
VR5_TRIF_CTAS=$best timeout 400 ncu --set full --clock-control none --import-source on -k regex:tri_kernel -s 30 -c 2 -f -o gpurun_out/trif_best \
  python bench.py --steps 12 --warmup 3 --no-cpu-baseline --sampler trilinear-fma > gpurun_out/trif_ncu.log 2>&1
echo "ncu rc=$? (VR5_TRIF_CTAS=$best)"; echo $best > gpurun_out/trif_best_ctas.txt
