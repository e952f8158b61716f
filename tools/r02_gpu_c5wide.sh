#!/bin/bash
# kernel variant 10 (wide window, 2048^3): parity on the forced small cases, the C5 full-size test, then the c5 record with it and without
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "wide_window" 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_full_size.py -m gpu -x -q -k "c5" -s 2>&1 | tail -4
(timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-policies > gpurun_out/r02_c5_wide.json 2> gpurun_out/r02_c5_wide.err; echo "c5 wide rc=$?"); tail -2 gpurun_out/r02_c5_wide.err
(VR5_P2_WIDE=-1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-policies > gpurun_out/r02_c5_tex.json 2> gpurun_out/r02_c5_tex.err; echo "c5 tex rc=$?"); tail -2 gpurun_out/r02_c5_tex.err
python - <<'PY'
import json
for n in ("r02_c5_wide","r02_c5_tex"):
    try:
        d=json.load(open("gpurun_out/%s.json"%n))
        c=d["c5"]
        print(n, "c3 ms %.4f | c5 ms %.3f gsamples %.1f %s verify %s"%(d["ms_per_step"], c["ms_per_step"], c["gsamples"], c["kernel"], {k:v for k,v in c["verify"].items() if k not in("what","oracle_note")}))
    except Exception as e:
        print(n, "failed", e)
PY
