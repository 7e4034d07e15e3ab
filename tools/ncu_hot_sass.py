"""Top stall-sampled SASS instructions per kernel of an .ncu-rep (developer tool).
Usage: python tools/ncu_hot_sass.py report.ncu-rep [kernel-index] [top-n] [name-filter]"""
import csv, io, subprocess, sys
path = sys.argv[1]; want = int(sys.argv[2]) if len(sys.argv) > 2 else 0; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
kernels, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == 'Kernel Name':
        cur = {'name': row[1], 'rows': [], 'head': None}; kernels.append(cur)
    elif cur is not None and cur['head'] is None:
        cur['head'] = row
    elif cur is not None and row:
        cur['rows'].append(row)
if len(sys.argv) > 4:
    kernels = [k for k in kernels if sys.argv[4] in k['name']]
k = kernels[want]; h = {n: i for i, n in enumerate(k['head'])}
tot = sum(int(r[h['# Samples']]) for r in k['rows'])
print(k['name'][:100], 'total samples', tot, 'instructions', len(k['rows']))
stalls = [n for n in k['head'] if n.startswith('stall_') and 'Not Issued' not in n]
order = sorted(range(len(k['rows'])), key=lambda i: -int(k['rows'][i][h['# Samples']]))[:topn]
for i in sorted(order):
    r = k['rows'][i]; s = int(r[h['# Samples']])
    top = sorted(((int(r[h[n]]), n[6:]) for n in stalls), reverse=True)[:2]
    print('%5d %5.1f%% [%4d] %-70s %s' % (s, 100.0 * s / max(tot, 1), i, r[h['Source']].strip()[:70], ' '.join('%s:%d' % (n, c) for c, n in top if c)))
