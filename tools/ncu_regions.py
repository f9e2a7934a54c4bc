#!/usr/bin/env python
"""Per-region stall breakdown from `ncu --page source --csv --print-source sass` output.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv; python tools/ncu_regions.py src.csv a:b:name ..."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hs = [i for i, r in enumerate(rows) if 'Source' in r]
h = hs[0]
end = hs[1] if len(hs) > 1 else len(rows)
hdr = rows[h]
ii = hdr.index('Instructions Executed')
body = [r for r in rows[h + 1:end] if len(r) > ii]
cols = ['stall_barrier', 'stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_math', 'stall_wait', 'stall_not_selected',
        'stall_selected', 'stall_dispatch', 'stall_branch_resolving', 'stall_no_inst', 'stall_lg', 'stall_sleep']
idx = {c: hdr.index(c) for c in cols if c in hdr}
ci = [i for i, c in enumerate(hdr) if c.startswith('Warp Stall Sampling (All')][0]


def f(x):
    try:
        return float(x or 0)
    except ValueError:
        return 0.0


tot = sum(f(r[ci]) for r in body)
print("total samples", tot, "lines", len(body))
for spec in sys.argv[2:]:
    a, b, name = spec.split(':')
    a, b = int(a), int(b)
    s = sum(f(r[ci]) for r in body[a:b])
    ins = sum(f(r[ii]) for r in body[a:b])
    parts = {c: sum(f(r[i]) for r in body[a:b]) for c, i in idx.items()}
    top = sorted(parts.items(), key=lambda kv: -kv[1])[:6]
    print("%-10s %5d-%5d samples %6.0f (%4.1f%%) instr %6.2fM | %s" % (
        name, a, b, s, 100 * s / tot, ins / 1e6, "  ".join("%s %.0f" % (k.replace('stall_', ''), v) for k, v in top)))
