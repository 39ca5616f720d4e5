#!/bin/bash
# round-2 multi-GPU run: bash tools/gpu_nn.sh N -- both bench arms as the driver launches them (torchrun, one rank per GPU)
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --impl reference --gpus $N --steps 5 --warmup 2 > gpurun_out/r02_bench_${N}gpu_reference.json 2> gpurun_out/r02_bench_${N}gpu_reference.err; echo "ref rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err; echo "bench rc=$?"; tail -4 gpurun_out/r02_bench_${N}gpu.err
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_${N}gpu.json')); r=json.load(open('gpurun_out/r02_bench_${N}gpu_reference.json'))
print('value', d['value'], 'burst', d['burst']['value'], 'e2e', d['e2e']['value'], d['e2e']['h2d_gbs_per_rank'])
print('check', d['check']['mismatches_vs_port_per_rank'], d['check']['frames_read_ok_per_rank'], 'sharded', d['sharded_check']['mismatches_vs_port_per_rank'])
print('ingest', d['ingest']['value'], d['ingest']['h2d_gbs_per_rank'], d['ingest']['golden_line_mismatches_per_rank'])
print('config3', d['config3']['value'], d['config3']['frames_per_gpu'], d['config3']['check']['mismatches_vs_port_per_rank'], d['config3']['e2e']['value'])
print('set2', d['set2']['value'], d['set2']['check']['mismatches_vs_port_per_rank']); print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], d['cpu_baseline']['mismatches_vs_gpu'], 'ref arm', r['value'], r['cpu_baseline']['cores'])"
nvidia-smi topo -m | head -12 | cut -c1-160; nproc; free -g | head -2
