set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v12.log 2>&1; tail -3 $O/pytest_gpu_v12.log
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 32768 --json $O/jpeg_v12.json > $O/jpeg_v12.log 2>&1; grep -A3 '"ms_per_batch"' $O/jpeg_v12.log; grep melf_read_jpeg $O/jpeg_v12.log | tail -2
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode --json $O/jpeg_v12_64k.json > $O/jpeg_v12_64k.log 2>&1; grep -A3 '"ms_per_batch"' $O/jpeg_v12_64k.log; grep melf_read_jpeg $O/jpeg_v12_64k.log | tail -2
MELF_JPEG_HOST_THREADS=1 MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode > $O/jpeg_v12_64k_t1.log 2>&1; grep -A3 '"ms_per_batch"' $O/jpeg_v12_64k_t1.log; grep melf_read_jpeg $O/jpeg_v12_64k_t1.log | tail -1
nproc
