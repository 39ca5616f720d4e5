set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v12.log 2>&1; tail -3 $O/pytest_gpu_v12.log
rm -f $O/v12_sweep.txt
for N in 65536 131072; do for L in 16 32; do
  T=n${N}_l${L}
  MELF_JPEG_TRACE=1 MELF_JPEG_LANES=$L python tools/bench_jpeg.py --streams $N --reps 3 --skip-decode --json $O/sweep12_$T.json > $O/sweep12_$T.log 2>&1
  echo "$T $(python -c "import json;print(round(json.load(open('$O/sweep12_$T.json'))['e2e']['frames_per_s']))") $(grep melf_read_jpeg_batch $O/sweep12_$T.log | tail -1)" >> $O/v12_sweep.txt
done; done
cat $O/v12_sweep.txt
