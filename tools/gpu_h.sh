#!/bin/bash
# round-2 run H: staging variants A/B (default build vs the straight global loads)
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_h.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu_h.log
for v in nofence fence; do
  if [ $v = fence ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_narrow/libmeterelf_b200.so; fi
  python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_j_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_j_$v.txt
  python tools/run_locate.py 6 | tail -1
  MELF_TC_DEBUG_SKIP=8 python tools/tc_timeline.py gpurun_out/tl_$v.raw 2>&1 | tail -2
  python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --api-files 0 --min-seconds 0.5 > gpurun_out/r02_bench_j_$v.json 2> gpurun_out/r02_bench_j_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_j_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['check']['mismatches_vs_port_per_rank'], d['set2']['kernel_ms_per_step'], d['set2']['check']['mismatches_vs_port_per_rank'])"
done
