#!/bin/bash
# round-2 run D2: ingest batch-size sweep, default and 32 hardware queues
for n in 65536 131072; do
  for c in 8 32; do
    echo "== n=$n connections=$c"
    CUDA_DEVICE_MAX_CONNECTIONS=$c MELF_JPEG_TRACE=1 timeout 400 python tools/ingest_stages.py $n 2>&1 | tail -2 | cut -c1-330
  done
done
