#!/bin/bash
# the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_noalloc.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu_noalloc.log
