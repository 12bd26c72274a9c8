#!/usr/bin/env python
"""Summarise ncu output brought back from the GPU box.

    python scripts/ncu_summary.py launches gpurun_out/x_launches.csv      # per-kernel time shares
    python scripts/ncu_summary.py raw gpurun_out/x_full_raw.csv           # key metrics of `--page raw --csv`
"""
import collections
import csv
import io
import sys

KEYS = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct', 'sm__inst_executed.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_elapsed.max']


def launches(path):
    lines = [l for l in open(path).read().splitlines() if not l.startswith('==')]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(io.StringIO("\n".join(lines))):
        if row.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        k = row['Kernel Name'].split('(')[0]
        v = float(row['Metric Value'].replace(',', ''))
        v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'usecond': 1.0, 'nsecond': 1e-3, 'msecond': 1e3}.get(row['Metric Unit'], 1.0)
        tot[k] += v
        cnt[k] += 1
    s = sum(tot.values())
    print(f"{'kernel':40s} {'n':>5s} {'avg us':>10s} {'share':>7s}")
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        print(f"{k:40s} {cnt[k]:5d} {v / cnt[k]:10.2f} {100 * v / s:6.1f}%")


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    for d in data:
        print('==', d[idx['Kernel Name']][:60])
        for w in KEYS:
            if w in idx:
                print(f"    {w:70s} {d[idx[w]]:>16s} {units[idx[w]]}")


if __name__ == "__main__":
    {'launches': launches, 'raw': raw}[sys.argv[1]](sys.argv[2])
