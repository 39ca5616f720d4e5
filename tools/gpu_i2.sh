#!/bin/bash
# round-2 run I2: what bounds a lone Huffman launch?  debug builds: 1 = all lanes read one stream, 2 = all stores to the spare block, 3 = both
for d in 0 1 2 3; do
  echo "== debug=$d"; MELF_LIB=$PWD/meterelf_b200/_lib_hd$d/libmeterelf_b200.so MELF_JPEG_SLOTS=1 timeout 300 python tools/ingest_stages.py 8192 2>&1 | tail -1 | cut -c90-260
done
