#!/bin/bash
# round-2 run F (1 GPU): full GPU test suite; ncu source-level capture of the dial kernel
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_f.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02_pytest_gpu_f.log
python tools/run_path.py 4 > gpurun_out/plain_f.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:melf_dials_kernel -s 2 -c 1 -o gpurun_out/r02_dials python tools/run_path.py 4 > gpurun_out/ncu_f.log 2>&1
tail -2 gpurun_out/plain_f.log gpurun_out/ncu_f.log
