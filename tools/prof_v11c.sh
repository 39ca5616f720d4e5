# v11 final ingest numbers (run under gpurun): parity, ingest bench at the tuned defaults, launch list + Huffman capture
set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v11.log 2>&1; tail -3 $O/pytest_gpu_v11.log
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 32768 --json $O/jpeg_v11.json > $O/jpeg_v11.log 2>&1; tail -12 $O/jpeg_v11.log
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode --json $O/jpeg_v11_64k.json > $O/jpeg_v11_64k.log 2>&1; tail -12 $O/jpeg_v11_64k.log
python bench.py > $O/bench_v11.json 2> $O/bench_v11.err
CMD="python tools/bench_jpeg.py --streams 32768 --reps 1 --skip-decode"
$CMD > $O/j2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf -c 300 --csv --log-file $O/v11_jpeg_launches.csv $CMD > $O/ncu1.log 2>&1
$CMD > $O/j3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:melf_jpeg_huffman -s 2 -c 1 -o $O/v11_jpeg_huffman -f $CMD > $O/ncu2.log 2>&1
ls $O | head -50
