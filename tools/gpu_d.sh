#!/bin/bash
# round-2 run D: where the locate kernel's time goes: skip experiments (timeline) + ncu with source for two staging variants
for skip in 0 1 2 4 3 7 8 15; do
  echo "== MELF_TC_DEBUG_SKIP=$skip (1 no staging, 2 no window sums, 4 no drain, 8 no MMAs)"
  MELF_TC_DEBUG_SKIP=$skip python tools/tc_timeline.py gpurun_out/tl_skip$skip.raw 2>&1 | tail -3
  MELF_TC_DEBUG_SKIP=$skip python tools/run_locate.py 5 2>&1 | tail -1
done > gpurun_out/r02_skip_experiments_d.txt 2>&1
cat gpurun_out/r02_skip_experiments_d.txt
python tools/run_locate.py 4 > gpurun_out/plain_d.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:melf_match_fused_ov -s 2 -c 1 -o gpurun_out/r02_locate_straight python tools/run_locate.py 4 > gpurun_out/ncu_d1.log 2>&1
tail -3 gpurun_out/ncu_d1.log
