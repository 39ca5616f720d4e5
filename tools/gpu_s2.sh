#!/bin/bash
# round-2 run S2: warps per Huffman CTA at small batches
for N in 8192 16384; do for w in 8 1 4 8 1 4; do
  echo "== cta_warps=$w n=$N"; MELF_JPEG_CTA_WARPS=$w timeout 100 python tools/ingest_stages.py $N 2>&1 | tail -1 | cut -c1-75
done; done
