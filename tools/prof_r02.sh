# round-2 profile run (under gpurun): launch list of the bench command, full ncu captures of the locate and dial kernels
set -x
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --min-seconds 0 --check-frames 32 --cpu-frames-per-core 4 --e2e-frames 256 --ingest-streams 8192 --config3-frames 2048 --api-files 0"
$CMD > $O/r02_prof_plain.json 2> $O/r02_prof_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf -c 600 --csv --log-file $O/r02_launches.csv $CMD > $O/ncu_p1.log 2>&1
python tools/ncu_summary.py launches $O/r02_launches.csv > $O/r02_launches.txt; cat $O/r02_launches.txt
python tools/run_path.py 5 > $O/r02_prof_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:melf_(match_fused_ov|dials)_kernel" -s 6 -c 2 -f -o $O/r02_final python tools/run_path.py 5 > $O/ncu_p2.log 2>&1
tail -2 $O/ncu_p2.log
