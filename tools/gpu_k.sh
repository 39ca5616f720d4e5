#!/bin/bash
# round-2 run K: full tests, full bench (TMA staging, sorted Huffman lanes), ingest A/B without the sort
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_k.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_k.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_k.json 2> gpurun_out/r02_bench_k.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_k.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_k.json'))
print('value', d['value'], 'burst', d['burst']['value'], 'e2e', d['e2e']['value'], d['kernel_ms_per_step'], d['clocks'])
print('ingest', d['ingest']['value'], d['ingest']['golden_line_mismatches_per_rank'], 'api', d.get('api'))
print('config3', d['config3']['value'], d['config3']['burst']['value'], 'set2', d['set2']['value'])
print('cpu', d['cpu_baseline'])"
MELF_JPEG_SORT=0 python bench.py --steps 5 --warmup 3 --config3-frames 0 --micro-bench 0 --api-files 0 --min-seconds 0.2 > gpurun_out/r02_bench_k_nosort.json 2> gpurun_out/r02_bench_k_nosort.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_k_nosort.json')); print('ingest without the sort', d['ingest']['value'], d['ingest']['golden_line_mismatches_per_rank'])"
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r02_bench_k_reference.json 2> gpurun_out/r02_bench_k_reference.err; echo "ref rc=$?"
