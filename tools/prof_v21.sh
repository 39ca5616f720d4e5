# v21 validation (run under gpurun)
set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q > $O/pytest_gpu_v21.log 2>&1; tail -4 $O/pytest_gpu_v21.log
for R in 1 2 3; do python tools/bench_jpeg.py --streams 65536 --reps 6 --skip-decode --json $O/jpeg_v21_64k_$R.json > $O/jpeg_v21_64k.log 2>&1; grep -A1 '"ms_per_batch"' $O/jpeg_v21_64k.log; done
python tools/bench_jpeg.py --streams 32768 --reps 5 --json $O/jpeg_v21.json > $O/jpeg_v21.log 2>&1; grep -E "frames_per_s" $O/jpeg_v21.log
python bench.py > $O/bench_v21.json 2> $O/bench_v21.err; wc -l $O/bench_v21.json
python -c "
import json;d=json.load(open('$O/bench_v21.json'));print(d['value'],d['e2e']['value'],d['roofline']['frac'],d['clocks']['samples'],d['ingest']['value'],d['ingest']['golden_line_mismatches'])"
