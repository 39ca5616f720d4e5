#!/bin/bash
# round-2 run J2: per-stream clocks / iterations of a lone lean Huffman launch (debug build 4)
MELF_LIB=$PWD/meterelf_b200/_lib_hd4/libmeterelf_b200.so MELF_HUFF_DUMP=gpurun_out/huff_dump.txt MELF_JPEG_SLOTS=1 timeout 300 python tools/ingest_stages.py 8192 2>&1 | tail -1 | cut -c90-200
