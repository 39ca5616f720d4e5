# round-1 v9 measurement script (run under gpurun): parity tests, bench lines, ncu launch list and full captures
set -x
timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_v9.log 2>&1; tail -2 gpurun_out/pytest_gpu_v9.log
python bench.py > gpurun_out/bench_v9.json 2> gpurun_out/bench_v9.err || exit 1
python bench.py --size 1280x720 --frames 4096 --steps 20 > gpurun_out/bench_v9_1280x720.json 2>> gpurun_out/bench_v9.err
CMD="python bench.py --steps 2 --warmup 1 --cpu-frames 64 --e2e-frames 256"
$CMD > gpurun_out/b2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf -c 400 --csv --log-file gpurun_out/v9_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/b3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_match_fused_ov -s 1 -c 1 -o gpurun_out/v9_fused -f $CMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
$CMD > gpurun_out/b4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_dials -s 1 -c 1 -o gpurun_out/v9_dials -f $CMD > gpurun_out/ncu3.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_v9_reference.json 2>> gpurun_out/bench_v9.err
MELF_TC_TIMELINE=x python tools/tc_timeline.py > /dev/null 2>&1; cp gpurun_out/timeline.txt gpurun_out/v9_timeline.txt
ls gpurun_out
