#!/bin/bash
# round-2 run X: tensor-copy staging with / without the L2 prefetch of the next pass's crop rows
for v in base l2pf; do
  if [ $v = l2pf ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_narrow/libmeterelf_b200.so; fi
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "locate or match or synthetic or golden" 2>&1 | tail -1
  timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_x_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_x_$v.txt
  timeout 100 python tools/run_locate.py 6 | tail -1
  timeout 300 python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --api-files 0 --min-seconds 0.5 > gpurun_out/r02_bench_x_$v.json 2> gpurun_out/r02_bench_x_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_x_$v.json')); print('$v', d['value'], d['burst']['value'], d['kernel_ms_per_step'], d['clocks']['sm_mhz'], d['check']['mismatches_vs_port_per_rank'], d['set2']['kernel_ms_per_step'], d['set2']['check']['mismatches_vs_port_per_rank'])"
done
