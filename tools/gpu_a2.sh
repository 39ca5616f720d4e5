#!/bin/bash
# round-2 run A2: all-zero Toeplitz K-steps skipped (default) vs all eight per pair (MELF_TC_KPP8=1)
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_a2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_a2.log
for v in kpp7 kpp8; do
  if [ $v = kpp8 ]; then export MELF_TC_KPP8=1; fi
  timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_a2_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_a2_$v.txt
  timeout 100 python tools/run_locate.py 8 | tail -1
  MELF_TC_DEBUG_SKIP=7 timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw 2>&1 | tail -3 | head -2
  timeout 300 python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --api-files 0 --min-seconds 0.5 > gpurun_out/r02_bench_a2_$v.json 2> gpurun_out/r02_bench_a2_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_a2_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['check']['mismatches_vs_port_per_rank'], d['set2']['kernel_ms_per_step'], d['set2']['check']['mismatches_vs_port_per_rank'], d['roofline']['frac'])"
done
