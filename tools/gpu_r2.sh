#!/bin/bash
# round-2 run R2: lean Huffman loop in 4-warp CTAs sharing one table set (default) vs 1-warp CTAs (MELF_JPEG_CTA_WARPS=1)
timeout 600 python -m pytest tests/test_jpeg_gpu.py -m gpu -x -q 2>&1 | tail -2
for N in 32768 65536 131072; do for w in 4 1; do
  echo "== cta_warps=$w n=$N"; MELF_JPEG_CTA_WARPS=$w timeout 300 python tools/ingest_stages.py $N 2>&1 | tail -1 | cut -c1-75
done; done
echo "== cta_warps=4 n=65536"; MELF_JPEG_CTA_WARPS=4 timeout 300 python tools/ingest_stages.py 65536 2>&1 | tail -1 | cut -c1-75
echo "== cta_warps=1 n=65536"; MELF_JPEG_CTA_WARPS=1 timeout 300 python tools/ingest_stages.py 65536 2>&1 | tail -1 | cut -c1-75
