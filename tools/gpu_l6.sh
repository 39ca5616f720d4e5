#!/bin/bash
# round-2 run L6: lanes per warp x hardware queues with the lean loop and the bucketed order
run() { echo "== $*"; env "$@" timeout 300 python tools/ingest_stages.py ${N:-32768} 2>&1 | tail -1 | cut -c1-75; }
for N in 32768 65536; do export N
run A=1
run MELF_JPEG_LANES=32
run CUDA_DEVICE_MAX_CONNECTIONS=32
run CUDA_DEVICE_MAX_CONNECTIONS=32 MELF_JPEG_LANES=32
run CUDA_DEVICE_MAX_CONNECTIONS=16
done
