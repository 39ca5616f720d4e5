# final round-1 measurement (run under gpurun): parity, bench lines, ingest bench, launch lists
set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v19.log 2>&1; tail -3 $O/pytest_gpu_v19.log
python bench.py > $O/bench_v19.json 2> $O/bench_v19.err; wc -l $O/bench_v19.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_v19_reference.json 2>> $O/bench_v19.err
python bench.py --size 1280x720 --frames 4096 --steps 20 > $O/bench_v19_1280x720.json 2>> $O/bench_v19.err
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 32768 --reps 5 --json $O/jpeg_v19.json > $O/jpeg_v19.log 2>&1; grep -E "frames_per_s|done at" $O/jpeg_v19.log | tail -4
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 65536 --reps 6 --skip-decode --json $O/jpeg_v19_64k.json > $O/jpeg_v19_64k.log 2>&1; grep -E "frames_per_s|done at" $O/jpeg_v19_64k.log | tail -2
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 131072 --reps 4 --skip-decode --json $O/jpeg_v19_128k.json > $O/jpeg_v19_128k.log 2>&1; grep -E "frames_per_s|done at" $O/jpeg_v19_128k.log | tail -2
CMD="python bench.py --steps 2 --warmup 1 --cpu-frames 64 --e2e-frames 256 --ingest-streams 8192"
$CMD > $O/b2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf -c 400 --csv --log-file $O/v19_launches.csv $CMD > $O/ncu1.log 2>&1
python tools/ncu_summary.py launches $O/v19_launches.csv
CMD="python tools/bench_jpeg.py --streams 32768 --reps 1 --skip-decode"
$CMD > $O/j3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_jpeg_huffman -s 2 -c 1 -o $O/v19_jpeg_huffman -f $CMD > $O/ncu2.log 2>&1
