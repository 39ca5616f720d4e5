#!/usr/bin/env python
"""Condense ncu output for profiles/: `ncu_summary.py launches <csv>` groups a
gpu__time_duration launch list by kernel; `ncu_summary.py rep <file.ncu-rep>` prints the
handful of raw-page metrics the roofline numbers are built from."""
import csv
import subprocess
import sys
from collections import OrderedDict

KEEP = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__warps_active.avg.pct_of_peak_sustained_active',
    'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active',
    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'lts__t_bytes.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
    'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
    'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
    'launch__shared_mem_per_block_dynamic', 'smsp__inst_executed.sum',
    'sm__cycles_elapsed.avg', 'sm__cycles_active.avg',
]


def launches(path):
    with open(path) as fp:
        lines = [l for l in fp if l.startswith('"')]
    groups = OrderedDict()
    for r in csv.DictReader(lines):
        key = (r['Kernel Name'].split('(')[0], r['Grid Size'], r['Block Size'])
        groups.setdefault(key, []).append(float(r['Metric Value']) / 1e3)
    total = sum(sum(v) for v in groups.values())
    print('%-34s %-18s %-14s %5s %10s %10s %7s' % ('kernel', 'grid', 'block', 'n', 'avg_us', 'sum_us', 'share'))
    for ((name, grid, block), v) in groups.items():
        print('%-34s %-18s %-14s %5d %10.1f %10.1f %6.1f%%' % (
            name, grid, block, len(v), sum(v) / len(v), sum(v), 100 * sum(v) / total))


def rep(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], stdout=subprocess.PIPE,
                         stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    (hdr, units) = (rows[0], rows[1])
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        print('== %s  grid %s block %s' % (d.get('Kernel Name'), d.get('Grid Size'), d.get('Block Size')))
        for (h, u, v) in zip(hdr, units, vals):
            if h in KEEP:
                print('  %-78s %12s %s' % (h, v, u))


def traffic(*paths):
    """JSON {kernel: dram bytes read + written per launch} from full-set reports, the
    `roofline.traffic` values bench.py reads from profiles/r01_traffic.json."""
    import json
    import re
    scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    out = {}
    for path in paths:
        txt = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL, text=True).stdout
        rows = list(csv.reader(txt.splitlines()))
        (hdr, units) = (rows[0], rows[1])
        for vals in rows[2:]:
            d = dict(zip(hdr, zip(units, vals)))
            name = re.sub(r'^void ', '', d['Kernel Name'][1]).split('<')[0].split('(')[0]
            tot = 0.0
            for key in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
                (u, v) = d[key]
                tot += float(v.replace(',', '')) * scale[u]
            out[name] = int(tot)
            out[name + ':grid'] = d['Grid Size'][1]
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    {'launches': launches, 'rep': rep, 'traffic': traffic}[sys.argv[1]](*sys.argv[2:])
