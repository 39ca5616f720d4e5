set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v17.log 2>&1; tail -3 $O/pytest_gpu_v17.log
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 32768 --json $O/jpeg_v17.json > $O/jpeg_v17.log 2>&1; grep -B2 -A22 '"wall_ms_per_batch"' $O/jpeg_v17.log | grep -E "huffman|frames_per_s"; grep -A3 '"ms_per_batch"' $O/jpeg_v17.log; grep melf_read_jpeg $O/jpeg_v17.log | tail -2
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode --json $O/jpeg_v17_64k.json > $O/jpeg_v17_64k.log 2>&1; grep -A3 '"ms_per_batch"' $O/jpeg_v17_64k.log; grep melf_read_jpeg $O/jpeg_v17_64k.log | tail -2
for L in 8 32; do MELF_JPEG_LANES=$L python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode > $O/jpeg_v17_64k_l$L.log 2>&1; echo lanes $L; grep -A1 '"ms_per_batch"' $O/jpeg_v17_64k_l$L.log; done
CMD="python tools/bench_jpeg.py --streams 32768 --reps 1 --skip-decode"
$CMD > $O/j2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:melf -c 300 --csv --log-file $O/v17_jpeg_launches.csv $CMD > $O/ncu1.log 2>&1
python tools/ncu_summary.py launches $O/v17_jpeg_launches.csv
