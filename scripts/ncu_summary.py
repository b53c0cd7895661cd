"""Per-kernel summary of an `ncu --page details --csv` export. usage: python scripts/ncu_summary.py <details.csv>"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]; ix = {k: i for i, k in enumerate(hdr)}
want = ['Duration', 'DRAM Throughput', 'L1/TEX Cache Throughput', 'L2 Cache Throughput', 'Issue Slots Busy', 'Executed Ipc Active',
        'Registers Per Thread', 'Achieved Occupancy', 'No Eligible', 'Eligible Warps Per Scheduler', 'Active Warps Per Scheduler',
        'Dynamic Shared Memory Per Block', 'Memory Throughput', 'Mem Busy', 'Max Bandwidth', 'Mem Pipes Busy',
        'Warp Cycles Per Issued Instruction', 'Compute (SM) Throughput', 'Avg. Active Threads Per Warp',
        'Avg. Not Predicated Off Threads Per Warp', 'Grid Size', 'Block Size', 'L2 Hit Rate', 'L1/TEX Hit Rate']
seen = {}
for r in rows[1:]:
    if len(r) <= ix['Metric Value']: continue
    k = (r[ix['ID']], r[ix['Kernel Name']][:60], r[ix['Grid Size']])
    if r[ix['Metric Name']] in want:
        seen.setdefault(k, {})[r[ix['Metric Name']]] = r[ix['Metric Value']] + ' ' + r[ix['Metric Unit']]
done = set()
for k, v in seen.items():
    if k[1] in done and '--all' not in sys.argv: continue
    done.add(k[1])
    print(k)
    print('    ' + ' | '.join(f"{a}: {b}" for a, b in v.items()))
