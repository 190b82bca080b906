"""Per-phase (between BAR.SYNC) stall-reason table from an `ncu --page source --csv` dump.

    python tools/ncu_phase_stalls.py SRC.csv [units]

Each row is one phase of the block program: share of warp samples, instructions, and where
the samples of that phase were spent (stall reasons, % of the phase's samples).  Barrier-wait
samples are attributed to the phase that ENDS at the barrier.
"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[hi]; data = rows[hi + 1:]
units = float(sys.argv[2]) if len(sys.argv) > 2 else None
ix = {h: i for i, h in enumerate(hdr)}
reasons = ['stall_barrier', 'stall_long_sb', 'stall_short_sb', 'stall_wait', 'stall_math', 'stall_mio',
           'stall_no_inst', 'stall_not_selected', 'stall_selected', 'stall_branch_resolving', 'stall_dispatch', 'stall_lg']


def f(r, k):
    try:
        return float(r[ix[k]])
    except Exception:
        return 0.0


segs = []
cur = {'s': 0.0, 'i': 0.0, 'fp64': 0.0, 'wf': 0.0, **{k: 0.0 for k in reasons}}
for r in data:
    src = r[ix['Source']]
    cur['s'] += f(r, '# Samples'); cur['i'] += f(r, 'Instructions Executed')
    if any(op in src for op in ('DFMA', 'DADD', 'DMUL')):
        cur['fp64'] += f(r, 'Instructions Executed')
    cur['wf'] += f(r, 'L1 Wavefronts Shared')
    for k in reasons:
        cur[k] += f(r, k)
    if 'BAR.SYNC' in src:
        # samples waiting at the barrier land on the instructions right after BAR.SYNC in SASS order;
        # keep it simple: close the segment here
        segs.append(cur)
        cur = {'s': 0.0, 'i': 0.0, 'fp64': 0.0, 'wf': 0.0, **{k: 0.0 for k in reasons}}
segs.append(cur)
tot = sum(s['s'] for s in segs); toti = sum(s['i'] for s in segs)
print('seg  samples%  inst%  inst/unit fp64/unit wf/unit | ' + ' '.join(k.replace('stall_', '')[:8].rjust(8) for k in reasons))
for n, s in enumerate(segs):
    if s['s'] == 0 and s['i'] == 0:
        continue
    per = (lambda v: v / units) if units else (lambda v: v / 1e6)
    line = '%3d  %7.1f %6.1f %9.1f %9.1f %7.1f | ' % (n, 100 * s['s'] / tot, 100 * s['i'] / toti, per(s['i']), per(s['fp64']), per(s['wf']))
    line += ' '.join(('%8.1f' % (100 * s[k] / s['s'] if s['s'] else 0.0)) for k in reasons)
    print(line)
print('total samples', tot, 'total warp inst', toti)
