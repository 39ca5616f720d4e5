# round-1 v10 measurement script (run under gpurun): parity tests incl. the JPEG ingest row, bench lines,
# ingest bench, ncu launch list of the ingest path and full captures of its two heaviest kernels
set -x
O=gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v10.log 2>&1; tail -2 $O/pytest_gpu_v10.log
python bench.py > $O/bench_v10.json 2> $O/bench_v10.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_v10_reference.json 2>> $O/bench_v10.err
python tools/bench_jpeg.py --streams 32768 --json $O/jpeg_v10.json > $O/jpeg_v10.log 2>&1; tail -30 $O/jpeg_v10.log
CMD="python tools/bench_jpeg.py --streams 4096 --reps 2"
$CMD > $O/j2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf -c 200 --csv --log-file $O/v10_jpeg_launches.csv $CMD > $O/ncu1.log 2>&1
$CMD > $O/j3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_jpeg_huffman -s 1 -c 1 -o $O/v10_jpeg_huffman -f $CMD > $O/ncu2.log 2>&1
$CMD > $O/j4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_jpeg_idct -s 1 -c 1 -o $O/v10_jpeg_idct -f $CMD > $O/ncu3.log 2>&1
$CMD > $O/j5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_jpeg_color -s 1 -c 1 -o $O/v10_jpeg_color -f $CMD > $O/ncu4.log 2>&1
ls -la $O
