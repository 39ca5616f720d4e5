#!/bin/bash
# round-2 run M2: uploads on their own stream (default) vs on the chunks' streams (MELF_JPEG_COPY_STREAM=0)
timeout 600 python -m pytest tests/test_jpeg_gpu.py -m gpu -x -q 2>&1 | tail -2
for c in 1 0; do for N in 32768 65536 131072; do
  echo "== copy_stream=$c n=$N"; MELF_JPEG_COPY_STREAM=$c timeout 300 python tools/ingest_stages.py $N 2>&1 | tail -1 | cut -c1-75
done; done
MELF_PROF_TRACE=gpurun_out/ingest_trace64b.txt timeout 300 python tools/ingest_stages.py 65536 2>&1 | tail -1 | cut -c1-75
