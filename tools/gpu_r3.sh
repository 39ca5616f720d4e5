#!/bin/bash
# round-2 run R3: warps per Huffman CTA
for N in 32768 65536 131072; do for w in 16 8 16 8; do
  echo "== cta_warps=$w n=$N"; MELF_JPEG_CTA_WARPS=$w timeout 300 python tools/ingest_stages.py $N 2>&1 | tail -1 | cut -c1-75
done; done
