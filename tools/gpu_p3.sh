#!/bin/bash
# round-2 run P3: dial kernel after the code-size cuts -- parity, then time
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --steps 20 --warmup 5 --config3-frames 0 --ingest-streams 0 --micro-bench 0 --api-files 0 --min-seconds 0.5 > gpurun_out/r02_bench_p3.json 2> gpurun_out/r02_bench_p3.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_p3.json')); print(d['value'], d['burst']['value'], d['burst']['kernel_ms_per_step'], d['kernel_ms_per_step'], d['check']['mismatches_vs_port_per_rank'], d['set2']['kernel_ms_per_step'], d['set2']['check']['mismatches_vs_port_per_rank'])"
