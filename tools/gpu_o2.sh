#!/bin/bash
# round-2 run O2: more, smaller chunks in flight (16 slots) so that the last chunk's Huffman launch starts on less data
run() { echo "== $*"; env "$@" MELF_JPEG_TRACE=1 timeout 300 python tools/ingest_stages.py ${N:-32768} 2>&1 | tail -2 | cut -c1-165 | sed 's/summed.*//'; }
for N in 32768 65536; do export N
run MELF_JPEG_SLOTS=8
run MELF_JPEG_SLOTS=16 MELF_JPEG_MIN_WARPS=128
run MELF_JPEG_SLOTS=16
run MELF_JPEG_SLOTS=12 MELF_JPEG_MIN_WARPS=128
done
