#!/bin/bash
# round-2 run E: GPU tests (overlay oracle, drop-in suite); L2 prefetch of the next pass's crop rows A/B
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_e.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02_pytest_gpu_e.log
for v in prefetch noprefetch; do
  if [ $v = noprefetch ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_narrow/libmeterelf_b200.so; fi
  python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_e_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_e_$v.txt
  python tools/run_locate.py 6 | tail -1
  python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --min-seconds 0.5 > gpurun_out/r02_bench_e_$v.json 2> gpurun_out/r02_bench_e_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_e_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['check']['mismatches_vs_port_per_rank'], d['set2']['kernel_ms_per_step'])"
done
