#!/bin/bash
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_t.json 2> gpurun_out/r02_bench_t.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_t.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_t.json'))
print('value', d['value'], 'burst', d['burst']['value'], d['burst']['kernel_ms_per_step'], d['kernel_ms_per_step'])
r=d['roofline']; print('frac', r['frac'], 'sustained', r['sustained'], r['frac_of_int8_issue_rate'])
print('set2', d['set2']['value'], d['set2']['roofline']['frac'], d['set2']['roofline']['kernel'][:30])"
