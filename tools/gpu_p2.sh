#!/bin/bash
# round-2 run P2: dial kernel code size (unroll factors) against its time
for v in u42 u21 u11; do
  export MELF_LIB=$PWD/meterelf_b200/_lib_$v/libmeterelf_b200.so
  timeout 300 python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --api-files 0 --min-seconds 0.5 > gpurun_out/r02_bench_p_$v.json 2> gpurun_out/r02_bench_p_$v.err; echo "bench $v rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r02_bench_p_$v.json')); print('$v', d['value'], d['burst']['value'], d['burst']['kernel_ms_per_step'], d['kernel_ms_per_step'], d['check']['mismatches_vs_port_per_rank'], d['set2']['burst']['kernel_ms_per_step'] if 'burst' in d['set2'] else d['set2']['kernel_ms_per_step'])"
done
