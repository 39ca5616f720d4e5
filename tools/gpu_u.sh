#!/bin/bash
python tools/ingest_stages.py 32768 2>&1 | tail -1
MELF_JPEG_SLOTS=1 python tools/ingest_stages.py 32768 2>&1 | tail -1
