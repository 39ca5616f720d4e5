#!/bin/bash
# round-2 run on 2 GPUs: the bench as the driver launches it (torchrun, one rank per GPU), both arms
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err; echo "bench rc=$?"; tail -5 gpurun_out/r02_bench_2gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_2gpu.json'))
print('value', d['value'], 'burst', d['burst']['value'], 'e2e', d['e2e'])
print('check', d['check']); print('sharded', d['sharded_check']); print('ingest', d['ingest']['value'], d['ingest']['h2d_gbs_per_rank'], d['ingest']['golden_line_mismatches_per_rank'])
print('config3', d['config3']['value'], d['config3']['check'], d['config3']['e2e']['value']); print('set2', d['set2']['value'], d['set2']['check']); print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['mismatches_vs_gpu'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r02_bench_2gpu_reference.json 2> gpurun_out/r02_bench_2gpu_reference.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/r02_bench_2gpu_reference.json
nvidia-smi topo -m | head -8; nproc
