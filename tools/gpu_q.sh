#!/bin/bash
# round-2 run Q: full tests + full bench with the small-map path chosen by the cost model
python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_q.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu_q.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_q.json 2> gpurun_out/r02_bench_q.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_q.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_q.json'))
print('value', d['value'], 'burst', d['burst']['value'], 'e2e', d['e2e']['value'], d['kernel_ms_per_step'])
print('ingest', d['ingest']['value'], 'api', d['api']['value'])
print('config3', d['config3']['value'], 'set2', d['set2']['value'], d['set2']['burst']['value'], d['set2']['kernel_ms_per_step'], d['set2']['check'])
for r in d['micro_bench']['rows']: print(r)"
MELF_MATCH_SMALL=1 MELF_SET=2 python tools/run_locate.py 6 | tail -1
