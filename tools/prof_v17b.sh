set -x
O=gpurun_out
rm -f $O/v17_sweep.txt
for N in 16384 32768 65536 131072; do for L in 16 32; do for S in 4 8; do
  T=n${N}_l${L}_s${S}
  MELF_JPEG_TRACE=1 MELF_JPEG_LANES=$L MELF_JPEG_SLOTS=$S python tools/bench_jpeg.py --streams $N --reps 3 --skip-decode --json $O/sweep17_$T.json > $O/sweep17_$T.log 2>&1
  echo "$T $(python -c "import json;print(round(json.load(open('$O/sweep17_$T.json'))['e2e']['frames_per_s']))") $(grep melf_read_jpeg_batch $O/sweep17_$T.log | tail -1)" >> $O/v17_sweep.txt
done; done; done
cat $O/v17_sweep.txt
