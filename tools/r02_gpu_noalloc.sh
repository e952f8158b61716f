#!/bin/bash
# the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (the file is collected last)
mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r02_noalloc_alone.log); tail -3 gpurun_out/r02_noalloc_alone.log
(timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_full.log 2>&1; echo "full rc=$?" >> gpurun_out/r02_pytest_full.log); tail -5 gpurun_out/r02_pytest_full.log
