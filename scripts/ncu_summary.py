#!/usr/bin/env python
"""Summarise ncu output brought back from the GPU box.

    python scripts/ncu_summary.py launches gpurun_out/x_launches.csv      # per-kernel time shares
    python scripts/ncu_summary.py raw gpurun_out/x_full_raw.csv           # key metrics of `--page raw --csv`
    python scripts/ncu_summary.py traffic gpurun_out/x_full_raw.csv out.json [source-name]   # dram bytes per launch and bench stage
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


STAGE_OF = {'march_kernel': 'march', 'density_fwd_kernel': 'density_fwd', 'integrate_fwd_kernel': 'integrate_fwd',
            'app_fwd_tc_kernel': 'app_fwd', 'app_fwd_kernel': 'app_fwd', 'prep_weights_kernel': 'shade_fwd',
            'shade_fused_fwd_kernel': 'shade_fwd', 'composite_kernel': 'composite', 'integrate_bwd_kernel': 'integrate_bwd',
            'layer3_bwd_kernel': 'shade_bwd_out', 'prep_weights_bwd_kernel': 'shade_bwd_dgrad',
            'shade_fused_dgrad_kernel': 'shade_bwd_dgrad', 'wgrad_fused_kernel': 'shade_wgrad',
            'wgrad_unpermute_kernel': 'shade_wgrad', 'app_bwd_vm_kernel': 'app_bwd', 'app_bwd_vm_tc_kernel': 'app_bwd', 'app_bwd_kernel': 'app_bwd',
            'density_bwd_vm_runs_kernel': 'density_bwd', 'density_bwd_vm_staged_kernel': 'density_bwd', 'density_bwd_kernel': 'density_bwd'}


def _bytes(val, unit):
    return float(val.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(unit, 1)


def traffic(path, out, source=None):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch, summed over the kernels of each bench stage
    (one training step captured with `ncu --set full`)."""
    import json
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    res = {}
    for d in data:
        name = d[idx['Kernel Name']].split('(')[0].split('<')[0].split('::')[-1].replace('void ', '').strip()
        st = STAGE_OF.get(name)
        if st is None:
            continue
        b = sum(_bytes(d[idx[m]], units[idx[m]]) for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
        e = res.setdefault(st, {'per_kernel': {}, 'source': source or path})
        e['per_kernel'].setdefault(name, []).append(b)
    for e in res.values():       # a kernel captured twice (window straddles two steps) counts once: mean per launch
        e['kernels'] = sorted(e['per_kernel'])
        e['dram_bytes_per_launch'] = int(sum(sum(v) / len(v) for v in e.pop('per_kernel').values()))
    json.dump(res, open(out, 'w'), indent=1, sort_keys=True)
    for k, e in sorted(res.items()):
        print(f"{k:20s} {e['dram_bytes_per_launch'] / 1e6:10.2f} MB  {','.join(e['kernels'])}")


if __name__ == "__main__":
    {'launches': launches, 'raw': raw, 'traffic': traffic}[sys.argv[1]](*sys.argv[2:])
