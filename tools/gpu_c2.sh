#!/bin/bash
# round-2 run C2: ingest with more hardware queues (CUDA_DEVICE_MAX_CONNECTIONS) and fewer Huffman lanes per warp
run() { echo "== $*"; env "$@" MELF_JPEG_TRACE=1 timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -2; }
run A=1
run CUDA_DEVICE_MAX_CONNECTIONS=32
run CUDA_DEVICE_MAX_CONNECTIONS=32 MELF_JPEG_LANES=8
run CUDA_DEVICE_MAX_CONNECTIONS=32 MELF_JPEG_LANES=32
run MELF_JPEG_LANES=8
CUDA_DEVICE_MAX_CONNECTIONS=32 MELF_PROF_TRACE=gpurun_out/ingest_trace_c32.txt timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -1
