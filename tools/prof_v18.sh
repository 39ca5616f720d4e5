set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v18.log 2>&1; tail -3 $O/pytest_gpu_v18.log
MELF_LIB=$PWD/meterelf_b200/_lib_look10/libmeterelf_b200.so timeout 600 python -m pytest tests/test_jpeg_gpu.py -m gpu -x -q > $O/pytest_gpu_v18_look10.log 2>&1; tail -3 $O/pytest_gpu_v18_look10.log
for R in 1 2; do
MELF_JPEG_TRACE=1 python tools/bench_jpeg.py --streams 65536 --reps 6 --skip-decode --json $O/jpeg_v18_64k_$R.json > $O/jpeg_v18_64k_$R.log 2>&1; echo look9 run $R; grep -A1 '"ms_per_batch"' $O/jpeg_v18_64k_$R.log
MELF_LIB=$PWD/meterelf_b200/_lib_look10/libmeterelf_b200.so python tools/bench_jpeg.py --streams 65536 --reps 6 --skip-decode --json $O/jpeg_v18_64k_look10_$R.json > $O/jpeg_v18_64k_look10_$R.log 2>&1; echo look10 run $R; grep -A1 '"ms_per_batch"' $O/jpeg_v18_64k_look10_$R.log
done
python tools/bench_jpeg.py --streams 32768 --reps 5 --json $O/jpeg_v18.json > $O/jpeg_v18.log 2>&1; grep -E "huffman|frames_per_s" $O/jpeg_v18.log
MELF_LIB=$PWD/meterelf_b200/_lib_look10/libmeterelf_b200.so python tools/bench_jpeg.py --streams 32768 --reps 5 --json $O/jpeg_v18_look10.json > $O/jpeg_v18_look10.log 2>&1; grep -E "huffman|frames_per_s" $O/jpeg_v18_look10.log
