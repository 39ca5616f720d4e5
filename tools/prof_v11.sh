# round-1 v11 measurement script (run under gpurun): ingest with MCU-rect compaction + early stop; lanes-per-warp sweep
set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v11.log 2>&1; tail -3 $O/pytest_gpu_v11.log
python tools/bench_jpeg.py --streams 32768 --json $O/jpeg_v11.json > $O/jpeg_v11.log 2>&1; tail -12 $O/jpeg_v11.log
for L in 4 8 16 32; do
  MELF_JPEG_LANES=$L python tools/bench_jpeg.py --streams 32768 --json $O/jpeg_v11_lanes$L.json > $O/jpeg_v11_lanes$L.log 2>&1
done
MELF_JPEG_LANES=16 python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode --json $O/jpeg_v11_64k_lanes16.json > $O/jpeg_v11_64k_lanes16.log 2>&1
MELF_JPEG_LANES=32 python tools/bench_jpeg.py --streams 65536 --reps 3 --skip-decode --json $O/jpeg_v11_64k_lanes32.json > $O/jpeg_v11_64k_lanes32.log 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/jpeg_v11*.json')):
    d = json.load(open(f))
    print(f, 'e2e %.0f frames/s' % d['e2e']['frames_per_s'], {k: round(v, 2) for (k, v) in d.get('e2e_stages_ms', {}).items()},
          'decode huffman ms', round(d.get('decode', {}).get('stages_ms', {}).get('jpeg_huffman', 0), 2))
PY
