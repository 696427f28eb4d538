#!/bin/bash
# r2q: the no-alloc entry point's test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2q_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/r2q_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2q_pytest.log
