#!/bin/bash
# round-2 run H2: Huffman variants when the lanes of a warp decode DIFFERENT files (MELF_JPEG_SORT=0: stream order i % 79)
for s in 0 1; do for f in 1 0; do
  echo "== sort=$s fast=$f"; MELF_JPEG_SORT=$s MELF_JPEG_FAST=$f timeout 300 python tools/ingest_stages.py 32768 2>&1 | tail -1 | cut -c1-260
done; done
