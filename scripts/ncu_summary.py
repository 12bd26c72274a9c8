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
            'app_fwd_ws_kernel': 'app_fwd', 'app_fwd_tc_kernel': 'app_fwd', 'app_fwd_kernel': 'app_fwd', 'prep_weights_kernel': 'shade_fwd',
            'shade_fused_fwd_kernel': 'shade_fwd', 'composite_kernel': 'composite', 'integrate_bwd_kernel': 'integrate_bwd',
            'layer3_bwd_kernel': 'shade_bwd_out', 'prep_weights_bwd_kernel': 'shade_bwd_dgrad',
            'shade_fused_dgrad_kernel': 'shade_bwd_dgrad', 'wgrad_fused_kernel': 'shade_wgrad',
            'wgrad_unpermute_kernel': 'shade_wgrad', 'app_bwd_vm_kernel': 'app_bwd', 'app_bwd_vm_tc_kernel': 'app_bwd', 'app_bwd_kernel': 'app_bwd',
            'density_bwd_vm_runs_kernel': 'density_bwd', 'density_bwd_vm_staged_kernel': 'density_bwd', 'density_bwd_kernel': 'density_bwd'}


def _bytes(val, unit):
    return float(val.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(unit, 1)


SECTORS = {'l1_load_bytes': 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',      # sectors looked up in L1 by global loads
           'l2_read_bytes': 'lts__t_sectors_srcunit_tex_op_read.sum',                # sectors L2 served to the SMs
           'l2_red_bytes': 'lts__t_sectors_srcunit_tex_op_red.sum',                  # sectors reduced in L2 (red.global)
           'l2_write_bytes': 'lts__t_sectors_srcunit_tex_op_write.sum'}
PCTS = {'l1tex_pct': 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts_pct': 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'dram_pct': 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'tensor_pct': 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm_pct': 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'warps_active_pct': 'sm__warps_active.avg.pct_of_peak_sustained_active'}


def traffic(path, out, source=None):
    """Per bench stage, from one training step captured with `ncu --set full`: dram__bytes_read + write per launch
    (bench.py's roofline.traffic), the on-chip traffic of the stage's dominant kernel (L1 load sectors, L2 read /
    reduction / write sectors, 32 B each), ncu's unit percentages and the duration under the profiler."""
    import json
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    num = lambda d, m: float(d[idx[m]].replace(',', '')) if m in idx and d[idx[m]] not in ('', 'n/a') else 0.0
    res = {}
    for d in data:
        name = d[idx['Kernel Name']].split('(')[0].split('<')[0].split('::')[-1].replace('void ', '').strip()
        st = STAGE_OF.get(name)
        if st is None:
            continue
        b = sum(_bytes(d[idx[m]], units[idx[m]]) for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
        rec = {'dram': b, 'us': num(d, 'gpu__time_duration.sum')}
        rec.update({k: num(d, m) * 32 for k, m in SECTORS.items()})
        rec.update({k: num(d, m) for k, m in PCTS.items()})
        rec['red_instructions'] = num(d, 'smsp__inst_executed_op_global_red.sum')
        e = res.setdefault(st, {'per_kernel': {}, 'source': source or path})
        e['per_kernel'].setdefault(name, []).append(rec)
    for e in res.values():       # a kernel captured twice (window straddles two steps) counts once: mean per launch
        pk = e.pop('per_kernel')
        mean = lambda v, key: sum(r[key] for r in v) / len(v)
        e['kernels'] = sorted(pk)
        e['dram_bytes_per_launch'] = int(sum(mean(v, 'dram') for v in pk.values()))
        top = max(pk, key=lambda k: mean(pk[k], 'us'))            # the stage's dominant kernel
        e['dominant_kernel'] = top
        e['us_under_ncu'] = round(mean(pk[top], 'us'), 2)
        for k in list(SECTORS) + ['red_instructions']:
            e[k] = int(mean(pk[top], k))
        for k in PCTS:
            e[k] = round(mean(pk[top], k), 1)
    json.dump(res, open(out, 'w'), indent=1, sort_keys=True)
    for k, e in sorted(res.items()):
        print(f"{k:18s} dram {e['dram_bytes_per_launch'] / 1e6:8.2f} MB  l1 {e['l1_load_bytes'] / 1e6:8.1f}  l2rd {e['l2_read_bytes'] / 1e6:8.1f}  "
              f"l2red {e['l2_red_bytes'] / 1e6:7.1f} MB | l1tex {e['l1tex_pct']:5.1f}% lts {e['lts_pct']:5.1f}% dram {e['dram_pct']:5.1f}% "
              f"tensor {e['tensor_pct']:5.1f}% | {e['us_under_ncu']:7.1f} us  {e['dominant_kernel']}")


if __name__ == "__main__":
    {'launches': launches, 'raw': raw, 'traffic': traffic}[sys.argv[1]](*sys.argv[2:])
