#!/bin/bash
# round-2 run Q2: full ncu capture of the dial kernel after the code-size cuts
python tools/run_path.py 5 > gpurun_out/q2_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:melf_dials_kernel" -s 3 -c 1 -f -o gpurun_out/r02_dials2 python tools/run_path.py 5 > gpurun_out/ncu_q2.log 2>&1
tail -2 gpurun_out/ncu_q2.log
