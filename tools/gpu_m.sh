#!/bin/bash
# round-2 run M: durations of the two kernels of the small-map path (ncu launch list, cold cache: shares, not absolutes)
MELF_MATCH_SMALL=1 MELF_SET=2 python tools/run_locate.py 3 > gpurun_out/plain_m.log 2>&1 &&
MELF_MATCH_SMALL=1 MELF_SET=2 ncu --metrics gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum --clock-control none -k regex:melf_small -c 4 --csv --log-file gpurun_out/r02_small_launches.csv python tools/run_locate.py 3 > gpurun_out/ncu_m.log 2>&1
cat gpurun_out/r02_small_launches.csv | cut -d, -f5,12- | tail -40
