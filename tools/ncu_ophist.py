"""Opcode histogram of an `ncu --page source --csv` dump, weighted by executed instructions and samples."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if 'Source' in r and 'Instructions Executed' in r:
        hi = i
        break
hdr = rows[hi]; data = rows[hi + 1:]
ix = {h: i for i, h in enumerate(hdr)}


def f(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0


tot = sum(f(r, 'Instructions Executed') for r in data)
ops = collections.Counter(); samp = collections.Counter()
for r in data:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[ix['Source']])
    op = m.group(2) if m else '?'
    ops[op] += f(r, 'Instructions Executed'); samp[op] += f(r, '# Samples')
print('total warp inst', tot)
ts = sum(samp.values())
for op, c in ops.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print('%-22s %6.2f%% inst  %6.2f%% samples' % (op, 100 * c / tot, 100 * samp[op] / ts))
