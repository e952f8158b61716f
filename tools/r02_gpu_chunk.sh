#!/bin/bash
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_full.log 2>&1; echo "full rc=$?" >> gpurun_out/r02_pytest_full.log); tail -3 gpurun_out/r02_pytest_full.log
for v in chunked single; do
  if [ $v = single ]; then export VR5_HOST_SINGLE=1; fi
  timeout 300 python bench.py --steps 20 --warmup 3 --no-c5 --no-policies --no-cpu-baseline > gpurun_out/chunk_$v.json 2> gpurun_out/chunk_$v.err
  timeout 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-c5 --no-policies --no-cpu-baseline > gpurun_out/chunk_c4_$v.json 2> gpurun_out/chunk_c4_$v.err
done
unset VR5_HOST_SINGLE
python - <<'PY'
import json
for n in ("chunk_chunked","chunk_single","chunk_c4_chunked","chunk_c4_single"):
    d=json.load(open("gpurun_out/%s.json"%n)); print(n, "blocking ms %.4f fps %.0f | headline %.4f e2e fps %.0f verify %s/%s"%(d["e2e_blocking"]["ms_per_frame"], d["e2e_blocking"]["fps"], d["ms_per_step"], d["e2e"]["fps"], d["verify"]["sequence_frames_equal"], d["verify"]["max_lsb"]))
PY
timeout 120 tools/volr_headless -n 256 -s 1920 1080 -tf semi -reps 5 | tail -3
