#!/bin/bash
# round-2 run Y: Toeplitz ring geometry (stages x K-steps per stage)
for v in n2k8 n4k4 n8k2 n3k4 n2k4; do
  if [ $v != n2k8 ]; then export MELF_LIB=$PWD/meterelf_b200/_lib_$v/libmeterelf_b200.so; fi
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "locate or match or synthetic or golden" 2>&1 | tail -1
  timeout 200 python tools/tc_timeline.py gpurun_out/tl_$v.raw > gpurun_out/r02_timeline_y_$v.txt 2>&1; tail -3 gpurun_out/r02_timeline_y_$v.txt
  timeout 100 python tools/run_locate.py 8 | tail -1
done
