#!/bin/bash
# round-2 run Z: what does a stage boundary cost the MMA stream?  (extra commit / fence / barrier poll per stage)
for v in base e1 e2 e3; do
  if [ $v != base ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_$v/libmeterelf_b200.so; fi
  timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_z_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_z_$v.txt
  timeout 100 python tools/run_locate.py 8 | tail -1
  MELF_TC_DEBUG_SKIP=7 timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw 2>&1 | tail -3 | head -2
  MELF_TC_DEBUG_SKIP=7 timeout 100 python tools/run_locate.py 8 | tail -1
done
