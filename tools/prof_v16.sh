set -x
O=gpurun_out
python bench.py > $O/bench_v16.json 2> $O/bench_v16.err; wc -l $O/bench_v16.json; tail -3 $O/bench_v16.err
python -c "
import json;d=json.load(open('$O/bench_v16.json'));print(d['value'],d['e2e']['value'],d['roofline']['frac'],d['clocks']);print(d['ingest'])"
