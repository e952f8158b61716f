#!/bin/bash
mkdir -p gpurun_out
python tools/dbg_ring.py 2>&1 | tail -10
VR5_WARP_FETCH=1 python tools/dbg_ring.py 2>&1 | grep "frames  7"
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02_noalloc_alone.log); tail -2 gpurun_out/r02_noalloc_alone.log
(timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_full.log 2>&1; echo "full rc=$?" >> gpurun_out/r02_pytest_full.log); tail -4 gpurun_out/r02_pytest_full.log
