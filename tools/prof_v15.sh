set -x
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_v15.log 2>&1; tail -3 $O/pytest_gpu_v15.log
python bench.py > $O/bench_v15.json 2> $O/bench_v15.err; wc -l $O/bench_v15.json; tail -3 $O/bench_v15.err
python -c "
import json;d=json.load(open('$O/bench_v15.json'));print(d['value'],d['e2e']['value'],d['roofline']['frac'],d['clocks'])"
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_v15_reference.json 2>> $O/bench_v15.err; wc -l $O/bench_v15_reference.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
