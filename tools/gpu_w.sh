#!/bin/bash
# round-2 run W: frame staging by tiled tensor copies (default) vs per-row bulk copies (MELF_STAGE_TMAP=0)
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_w.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_w.log
for v in tmap rows; do
  if [ $v = rows ]; then export MELF_STAGE_TMAP=0; fi
  timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_w_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_w_$v.txt
  timeout 100 python tools/run_locate.py 6 | tail -1
  timeout 300 python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --api-files 0 --min-seconds 0.5 > gpurun_out/r02_bench_w_$v.json 2> gpurun_out/r02_bench_w_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_w_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['check']['mismatches_vs_port_per_rank'], d['set2']['kernel_ms_per_step'], d['set2']['check']['mismatches_vs_port_per_rank'])"
done
