#!/bin/bash
# round-2 run L: the small-result-map path (set 2): parity test, timing against the Toeplitz kernel
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small_result_map" > gpurun_out/r02_pytest_gpu_l.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02_pytest_gpu_l.log
for v in small toeplitz; do
  if [ $v = toeplitz ]; then export MELF_MATCH_SMALL=0; else export MELF_MATCH_SMALL=1; fi
  MELF_SET=2 timeout 300 python tools/run_locate.py 6 | tail -1
  MELF_SET=2 timeout 300 python tools/run_path.py 6 | tail -1
done
