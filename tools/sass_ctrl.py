#!/usr/bin/env python3
"""SASS of one kernel with the scheduling control bits decoded (sm_70+ layout: high word bits 41.. = stall:4 yield:1 wbar:3 rbar:3
wait:6): which instruction sets which scoreboard and which instruction waits on it.  Finds false waits (a fast-path instruction
waiting on a scoreboard that a just-issued long-latency load shares).
   tools/sass_ctrl.py ngsxfem_b200/csrc/ngsx_rows.o 'k_rowsILi3ELi2E' [lo hi]      (SASS index range; default: summary)"""
import re, subprocess, sys, collections

obj, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(['cuobjdump', '-sass', obj], capture_output=True, text=True).stdout.splitlines()
ins, on = [], False
for k, line in enumerate(txt):
    if 'Function :' in line:
        on = pat in line
        continue
    if not on:
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/', line)
    if m:
        hi = re.search(r'/\* 0x([0-9a-f]{16}) \*/', txt[k + 1])
        c = int(hi.group(1), 16) >> 41
        ins.append(dict(addr=int(m.group(1), 16), text=' '.join(m.group(2).split()), stall=c & 15, yld=(c >> 4) & 1, wbar=(c >> 5) & 7,
                        rbar=(c >> 8) & 7, wait=(c >> 11) & 63))


def fmt(k, i):
    w = ''.join(str(b) for b in range(6) if i['wait'] >> b & 1)
    return f"{k:5d} {i['text'][:84]:84s} st{i['stall']:2d} {'W' + str(i['wbar']) if i['wbar'] != 7 else '  '} {'R' + str(i['rbar']) if i['rbar'] != 7 else '  '} {'wait ' + w if w else ''}"


if len(sys.argv) > 4:
    for k in range(int(sys.argv[3]), min(int(sys.argv[4]), len(ins))):
        print(fmt(k, ins[k]))
else:
    print(len(ins), 'instructions')
    ops = collections.Counter(i['text'].split()[1 if i['text'].startswith('@') else 0].split('.')[0] for i in ins)
    print(dict(ops.most_common(24)))
    for b in range(6):
        setters = [k for k, i in enumerate(ins) if i['wbar'] == b or i['rbar'] == b]
        waiters = [k for k, i in enumerate(ins) if i['wait'] >> b & 1]
        kinds = collections.Counter(ins[k]['text'].split()[1 if ins[k]['text'].startswith('@') else 0].split('.')[0] for k in setters)
        print(f'SB{b}: set by {len(setters)} ({dict(kinds)}), waited by {len(waiters)}')
