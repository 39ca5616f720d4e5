#!/bin/bash
# round-2 run R: ingest sweep with length-sorted lanes: lanes per warp, chunk size, slots
for lanes in 8 16 32; do MELF_JPEG_LANES=$lanes python tools/ingest_once.py 32768 2>&1 | tail -1; done
for chunk in 4096 16384; do MELF_JPEG_CHUNK=$chunk python tools/ingest_once.py 32768 2>&1 | tail -1; done
for slots in 4 8; do MELF_JPEG_SLOTS=$slots python tools/ingest_once.py 65536 2>&1 | tail -1; done
MELF_JPEG_LANES=32 MELF_JPEG_SLOTS=8 python tools/ingest_once.py 65536 2>&1 | tail -1
