#!/bin/bash
# round-2 run N3: lanes per warp once the uploads run back to back (Huffman stage throughput-bound)
for N in 32768 65536; do for l in 16 32 8; do
  echo "== lanes=$l n=$N"; MELF_JPEG_LANES=$l timeout 300 python tools/ingest_stages.py $N 2>&1 | tail -1 | cut -c1-75
done; done
