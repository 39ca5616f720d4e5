#!/bin/bash
# round-2 run F2: lean Huffman loop + redo pass (default) vs the careful loop (MELF_JPEG_FAST=0)
timeout 600 python -m pytest tests/test_jpeg_gpu.py -m gpu -x -q 2>&1 | tail -3
for f in 1 0; do
  echo "== fast=$f lone launch, 8192 streams"; MELF_JPEG_SLOTS=1 MELF_JPEG_FAST=$f timeout 300 python tools/ingest_stages.py 8192 2>&1 | tail -1 | cut -c1-250
  for s in 1 0; do
    echo "== fast=$f sort=$s n=32768"; MELF_JPEG_SORT=$s MELF_JPEG_FAST=$f timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -1 | cut -c1-250
  done
done
