#!/bin/bash
# round-2 run K2: order of the streams inside a chunk (MELF_JPEG_SORT: 0 stream order, 1 exact lengths, n = buckets of 2^n bytes)
for s in 10 1 0 8 12 10; do
  echo "== sort=$s"; MELF_JPEG_SORT=$s timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -1 | cut -c1-120
done
echo "== careful loop, sort=10"; MELF_JPEG_FAST=0 timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -1 | cut -c1-120
echo "== 65536, sort=10"; timeout 300 python tools/ingest_stages.py 65536 2>&1 | tail -1 | cut -c1-120
