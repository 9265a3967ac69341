#!/usr/bin/env python3
"""Hot SASS instructions of one kernel from an ncu report (stall samples per instruction).
   tools/ncu_hot.py gpurun_out/prof.ncu-rep k_rows [min_fraction] [lo hi]   (lo..hi: dump that SASS index range instead)"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
frac = float(sys.argv[3]) if len(sys.argv) > 3 else 0.008
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
start = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[start]
end = next((i for i in range(start + 1, len(rows)) if rows[i] and rows[i][0] == 'Address'), len(rows))   # first matching launch only
data = [r for r in rows[start + 1:end] if len(r) > 10]
col = {n: hdr.index(n) for n in ('# Samples', 'Instructions Executed', 'Source', 'stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_wait',
                                  'stall_lg', 'Avg. Threads Executed', 'stall_math', 'stall_barrier', 'stall_not_selected', 'stall_branch_resolving')}
S = lambda r, n: int(r[col[n]] or 0)
tot = sum(S(r, '# Samples') for r in data)
print('samples', tot, 'instructions', len(data), 'executed', sum(S(r, 'Instructions Executed') for r in data))
for n in ('stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_wait', 'stall_lg', 'stall_math', 'stall_barrier', 'stall_not_selected', 'stall_branch_resolving'):
    print(f'  {n}: {sum(S(r, n) for r in data)}')
if len(sys.argv) > 5:
    for k in range(int(sys.argv[4]), int(sys.argv[5])):
        r = data[k]
        print(k, r[col['Source']][:88].ljust(88), r[col['# Samples']], r[col['Instructions Executed']])
else:
    for k, r in enumerate(data):
        if S(r, '# Samples') > tot * frac:
            print(k, r[col['Source']][:70].ljust(70), S(r, '# Samples'), 'I', r[col['Instructions Executed']], 'thr', r[col['Avg. Threads Executed']],
                  'long', S(r, 'stall_long_sb'), 'short', S(r, 'stall_short_sb'), 'mio', S(r, 'stall_mio'), 'wait', S(r, 'stall_wait'))
