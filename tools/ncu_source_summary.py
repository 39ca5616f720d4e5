#!/usr/bin/env python
"""Per-instruction stall samples of one ncu capture condensed for profiles/:
    python tools/ncu_source_summary.py <file.ncu-rep> [top N]
prints the totals per stall reason, per opcode and the N SASS instructions with the most samples."""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    text = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], stdout=subprocess.PIPE,
                          stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(text.splitlines()))
    # one section per kernel: a "Kernel Name" row, a header row, then the instructions
    starts = [i for (i, r) in enumerate(rows) if r and r[0] == 'Kernel Name'] + [len(rows)]
    for (a, b) in zip(starts[:-1], starts[1:]):
        section(rows[a:b], top_n)
        print()


def section(rows, top_n):
    print('#', rows[0][1] if len(rows[0]) > 1 else rows[0])
    hdr = rows[1]
    data = [r for r in rows[2:] if len(r) >= len(hdr)]
    ix = {h: i for (i, h) in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    total = sum(int(r[ix['# Samples']] or 0) for r in data)
    executed = sum(int(r[ix['Instructions Executed']] or 0) for r in data)
    print('# warp stall samples: %d, warp instructions executed: %d' % (total, executed))
    per = collections.Counter()
    for r in data:
        for h in stalls:
            per[h] += int(r[ix[h]] or 0)
    print('\n## samples per stall reason')
    for (h, v) in per.most_common(10):
        print('%-26s %8d %5.1f %%' % (h, v, 100.0 * v / total))
    ops = collections.Counter()
    opn = collections.Counter()
    for r in data:
        toks = [t for t in r[ix['Source']].split() if not t.startswith('@')]
        if toks:
            ops[toks[0].split('.')[0]] += int(r[ix['# Samples']] or 0)
            opn[toks[0].split('.')[0]] += int(r[ix['Instructions Executed']] or 0)
    print('\n## per opcode: samples, share of samples, executed, share of executed')
    for (name, v) in ops.most_common(16):
        print('%-10s %8d %5.1f %%   %11d %5.1f %%' % (name, v, 100.0 * v / total, opn[name], 100.0 * opn[name] / executed))
    print('\n## the %d instructions with the most samples (in program order): index, SASS, samples, share, executed, main reasons' % top_n)
    top = sorted(range(len(data)), key=lambda i: -int(data[i][ix['# Samples']] or 0))[:top_n]
    for i in sorted(top):
        r = data[i]
        reasons = sorted(((h, int(r[ix[h]] or 0)) for h in stalls), key=lambda kv: -kv[1])[:2]
        print('%5d  %-64s %7s %5.1f %% %10s  %s' % (
            i, r[ix['Source']].strip()[:64], r[ix['# Samples']], 100.0 * int(r[ix['# Samples']]) / total,
            r[ix['Instructions Executed']], ', '.join('%s %d' % kv for kv in reasons if kv[1])))


if __name__ == '__main__':
    main()
