#!/bin/bash
# round-2 run B: new GPU tests; drain A/B (wide TMEM loads vs narrow) with the timeline and the short bench
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_b.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02_pytest_gpu_b.log
for v in wide narrow; do
  if [ $v = narrow ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_narrow/libmeterelf_b200.so; fi
  python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_b_$v.txt 2>&1; tail -12 gpurun_out/r02_timeline_b_$v.txt
  python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --min-seconds 0.5 > gpurun_out/r02_bench_b_$v.json 2> gpurun_out/r02_bench_b_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_b_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'])"
done
