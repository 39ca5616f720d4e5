#!/bin/bash
# round-2 run G2: full ncu capture (with source counters) of one Huffman launch, saturated (32768 streams in one chunk of 8192 x 1 slot is not possible: use 8192 streams, one slot)
export MELF_JPEG_SLOTS=1
true
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:melf_jpeg_huffman_kernel" -s 4 -c 1 -f -o gpurun_out/r02_huff python tools/ingest_stages.py 8192 > gpurun_out/ncu_huff.log 2>&1
tail -2 gpurun_out/ncu_huff.log
