"""Per-source-line instruction/stall attribution from `ncu --page source --csv --print-source cuda,sass`.
usage: python tools/ncu_lines.py <cs.csv> [kernel_index] [top]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
# the file is a sequence of (File Path, Function Name, header, rows...) sections; kernels repeat per profiled launch
sections, cur = [], None
for r in rows:
    if r and r[0] == 'File Path':
        cur = {'file': r[1], 'rows': []}; sections.append(cur)
    elif r and r[0] == 'Function Name' and cur is not None:
        cur['func'] = r[1]
    elif cur is not None:
        cur['rows'].append(r)
# group sections into launches: a new launch starts when a file repeats
launches, seen = [[]], set()
for s in sections:
    if s['file'] in seen:
        launches.append([]); seen = set()
    seen.add(s['file']); launches[-1].append(s)
L = launches[which]
tot_i = tot_s = 0
lines = collections.Counter(); samp = collections.Counter(); thr = collections.Counter(); text = {}
for s in L:
    hdr = s['rows'][0]
    ie, ws, te = hdr.index('Instructions Executed'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Thread Instructions Executed')
    for r in s['rows'][1:]:
        if len(r) <= te or not r[0]:
            continue            # sass rows have an empty Line No
        key = (s['file'].split('/')[-1], int(r[0]))
        try:
            i, w, t = int(r[ie] or 0), int(r[ws] or 0), int(r[te] or 0)
        except ValueError:
            continue
        lines[key] += i; samp[key] += w; thr[key] += t; text[key] = r[1].strip()[:110]
        tot_i += i; tot_s += w
print(f'launch {which}: {len(launches)} launches, warp instr {tot_i}, samples {tot_s}')
for key, c in lines.most_common(top):
    print(f'{100*c/tot_i:5.1f}% instr {100*samp[key]/max(1,tot_s):5.1f}% stall  thr/instr {thr[key]/max(1,c):5.1f}  {key[0]}:{key[1]}  {text[key]}')
