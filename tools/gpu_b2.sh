#!/bin/bash
# round-2 run B2: per-launch timeline of melf_read_jpeg_batch (where does the wall time go?)
MELF_PROF_TRACE=gpurun_out/ingest_trace.txt MELF_JPEG_TRACE=1 timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -6
