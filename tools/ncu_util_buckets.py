"""Bucket the warp instructions of one profiled launch by SIMD utilisation (threads per warp instruction)
from `ncu --page source --csv --print-source cuda,sass`: python tools/ncu_util_buckets.py <cs.csv> [launch]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
sections, cur = [], None
for r in rows:
    if r and r[0] == 'File Path':
        cur = {'file': r[1], 'rows': []}; sections.append(cur)
    elif r and r[0] == 'Function Name' and cur is not None:
        cur['func'] = r[1]
    elif cur is not None:
        cur['rows'].append(r)
launches, seen = [[]], set()
for s in sections:
    if s['file'] in seen:
        launches.append([]); seen = set()
    seen.add(s['file']); launches[-1].append(s)
buckets = collections.Counter(); thr = collections.Counter(); per_file = collections.Counter()
tot = 0
for s in launches[which]:
    hdr = s['rows'][0]
    ie, te = hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed')
    for r in s['rows'][1:]:
        if len(r) <= te or not r[0]:
            continue
        try:
            i, t = int(r[ie] or 0), int(r[te] or 0)
        except ValueError:
            continue
        if not i:
            continue
        u = t / i
        b = '<4' if u < 4 else '4-8' if u < 8 else '8-16' if u < 16 else '16-28' if u < 28 else '28-32'
        buckets[b] += i; thr[b] += t; tot += i
        per_file[(s['file'].split('/')[-1], b)] += i
print(f'launch {which}: warp instructions {tot}')
for b in ('<4', '4-8', '8-16', '16-28', '28-32'):
    print(f'  lanes/instr {b:6s}: {100 * buckets[b] / max(1, tot):5.1f}% of warp instructions, mean {thr[b] / max(1, buckets[b]):5.1f} lanes')
for (f, b), c in sorted(per_file.items(), key=lambda kv: -kv[1])[:12]:
    print(f'    {f:22s} {b:6s} {100 * c / tot:5.1f}%')
