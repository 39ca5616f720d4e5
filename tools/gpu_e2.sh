#!/bin/bash
# round-2 run E2: ingest chunk size x hardware queues (finer pipelining of upload -> Huffman)
run() { echo "== $*"; env "$@" MELF_JPEG_TRACE=1 timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -2 | cut -c1-260; }
run A=1
run MELF_JPEG_CHUNK=2048
run MELF_JPEG_CHUNK=2048 CUDA_DEVICE_MAX_CONNECTIONS=32
run MELF_JPEG_CHUNK=1024 CUDA_DEVICE_MAX_CONNECTIONS=32
run MELF_JPEG_CHUNK=2048 MELF_JPEG_LANES=16 CUDA_DEVICE_MAX_CONNECTIONS=32
run MELF_JPEG_CHUNK=2048 MELF_JPEG_LANES=4 CUDA_DEVICE_MAX_CONNECTIONS=32
