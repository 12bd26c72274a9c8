#!/usr/bin/env python
"""Top stalled SASS instructions of an `ncu --page source --csv` export.  python scripts/ncu_hot.py file.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
h = rows[1]
idx = {c: i for i, c in enumerate(h)}
data = rows[2:]


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return 0.0


tot = sum(num(r[idx['# Samples']]) for r in data)
print("total samples", tot, "instructions", len(data))
reasons = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
agg = {c: sum(num(r[idx[c]]) for r in data) for c in reasons}
print("by reason:", {k: int(v) for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
for r in sorted(data, key=lambda r: -num(r[idx['# Samples']]))[:N]:
    why = max(reasons, key=lambda c: num(r[idx[c]]))
    print(f"{num(r[idx['# Samples']]):7.0f} {100 * num(r[idx['# Samples']]) / tot:5.1f}% {why[6:]:>12s} ex={r[idx['Instructions Executed']]:>8s}  {r[idx['Source']][:100]}")
