#!/bin/bash
export MELF_MATCH_SMALL=1 MELF_SET=2
for v in a4b16_e256 a3b32_e128; do
  if [ $v = a3b32_e128 ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_narrow/libmeterelf_b200.so; fi
  echo "== $v"; timeout 300 python tools/run_locate.py 6 | tail -1
  python tools/run_locate.py 3 > gpurun_out/plain_m.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf_small -c 4 --csv --log-file gpurun_out/r02_small_$v.csv python tools/run_locate.py 3 > gpurun_out/ncu_m.log 2>&1
  grep gpu__time gpurun_out/r02_small_$v.csv | cut -d, -f5,15- | tail -4
done
