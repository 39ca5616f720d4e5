# v11 ingest sweep (run under gpurun): chunks in flight x streams per chunk x streams per warp, e2e leg only
set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v11.log 2>&1; tail -3 $O/pytest_gpu_v11.log
for N in 32768 65536; do
for S in 4 8; do for CH in 2048 4096 8192; do for L in 8 16; do
  T=n${N}_s${S}_c${CH}_l${L}
  MELF_JPEG_TRACE=1 MELF_JPEG_SLOTS=$S MELF_JPEG_CHUNK=$CH MELF_JPEG_LANES=$L python tools/bench_jpeg.py --streams $N --reps 3 --skip-decode --json $O/sweep_$T.json > $O/sweep_$T.log 2>&1
  echo "$T $(python -c "import json;print(round(json.load(open('$O/sweep_$T.json'))['e2e']['frames_per_s']))") $(grep melf_read_jpeg_batch $O/sweep_$T.log | tail -1)" >> $O/v11_sweep.txt
done; done; done; done
cat $O/v11_sweep.txt
