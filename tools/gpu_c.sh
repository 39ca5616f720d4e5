#!/bin/bash
# round-2 run C: GPU tests; staging A/B (cp.async ring in shared memory vs straight global loads)
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_c.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02_pytest_gpu_c.log
for v in ring straight; do
  if [ $v = straight ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_narrow/libmeterelf_b200.so; fi
  python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_c_$v.txt 2>&1; tail -4 gpurun_out/r02_timeline_c_$v.txt
  python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --min-seconds 0.5 > gpurun_out/r02_bench_c_$v.json 2> gpurun_out/r02_bench_c_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_c_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['check'], d['set2']['kernel_ms_per_step'])"
done
