"""Print the sampled hot region of one kernel from an ncu --page source --csv export.
usage: python scripts/ncu_hot.py <csv> [first_row last_row]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {k: i for i, k in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr) and r[ix['# Samples']] != '# Samples']
def S(r, c='# Samples'):
    try: return int(float(r[ix[c]] or 0))
    except ValueError: return 0
tot = sum(S(r) for r in data)
cols = ['stall_long_sb', 'stall_short_sb', 'stall_wait', 'stall_math', 'stall_not_selected', 'stall_dispatch', 'stall_mio', 'stall_branch_resolving', 'stall_selected', 'stall_barrier', 'stall_no_inst']
print('rows', len(data), 'samples', tot, {c: sum(S(r, c) for r in data) for c in cols})
best = sorted(range(len(data)), key=lambda i: -S(data[i]))[:8]
print('top rows', [(i, S(data[i])) for i in best])
if len(sys.argv) > 3:
    lo, hi = int(sys.argv[2]), int(sys.argv[3])
    for i in range(lo, min(hi, len(data))):
        r = data[i]
        st = ' '.join(f"{c.split('_', 1)[1][:6]}={r[ix[c]]}" for c in cols if r[ix[c]] not in ('0', ''))
        print(f"{i:4d} {S(r):6d} {r[ix['Instructions Executed']]:>9} {r[ix['Source']][:64]:64s} {st}")
