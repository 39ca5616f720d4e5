#!/bin/bash
# round-2 run G: what the staging phase is made of (loads vs conversion), with and without the MMAs
for skip in 0 16 32 8 24 40 6 22 38; do
  echo "== MELF_TC_DEBUG_SKIP=$skip (2 no window sums, 4 no drain, 8 no MMAs, 16 staging without loads, 32 staging without conversion)"
  MELF_TC_DEBUG_SKIP=$skip python tools/tc_timeline.py gpurun_out/tl_g$skip.raw 2>&1 | tail -2
  MELF_TC_DEBUG_SKIP=$skip python tools/run_locate.py 5 2>&1 | tail -1
done > gpurun_out/r02_skip_experiments_g.txt 2>&1
cat gpurun_out/r02_skip_experiments_g.txt
