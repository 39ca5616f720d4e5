#!/bin/bash
# round-2 final run, 1 GPU: GPU tests, both bench arms as the driver runs them, smoke
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_final.log
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_final_reference.json 2> gpurun_out/r02_bench_final_reference.err; echo "ref rc=$?"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_final.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_final.json')); r=json.load(open('gpurun_out/r02_bench_final_reference.json'))
print('value', d['value'], 'burst', d['burst']['value'], 'e2e', d['e2e']['value'], d['kernel_ms_per_step'], d['roofline']['frac'], d['roofline']['kernel'][:40])
print('ingest', d['ingest']['value'], 'api', d['api']['value'], 'config3', d['config3']['value'], 'set2', d['set2']['value'], d['set2']['roofline']['kernel'][:40])
print('reference arm', r['value'], r['cpu_baseline'])"
python -c "import __graft_entry__ as g; g.smoke()"
