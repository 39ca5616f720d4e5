#!/bin/bash
# round-2 run A: GPU tests, the new bench at N=1, the reference arm
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_a.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02_pytest_gpu_a.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_a.json 2> gpurun_out/r02_bench_a.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_a.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_ref_a.json 2> gpurun_out/r02_ref_a.err; echo "ref rc=$?"
nproc; free -g | head -2; nvidia-smi topo -m 2>/dev/null | head -12
