"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv) per kernel name (developer tool).
Usage: python tools/ncu_launch_summary.py launches.csv [first-launch-id] > profiles/....md"""
import csv
import re
import sys

path = sys.argv[1]
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = {}
total = n = 0
with open(path, newline='') as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get('Metric Name') != 'gpu__time_duration.sum' or int(r['ID']) < first:
        continue
    v = float(r['Metric Value'].replace(',', ''))
    unit = r['Metric Unit']
    ms = v * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 'nsecond': 1e-6, 'usecond': 1e-3, 'msecond': 1.0}.get(unit, 1e-6)
    name = r['Kernel Name']
    name = re.sub(r'^void ', '', name)
    name = re.sub(r'\(anonymous namespace\)::|<unnamed>::', '', name)
    name = name.split('(')[0][:80]
    a = rows.setdefault(name, [0.0, 0])
    a[0] += ms; a[1] += 1
    total += ms; n += 1
print('| share | ms | launches | kernel |\n|---|---|---|---|')
for name, (ms, c) in sorted(rows.items(), key=lambda kv: -kv[1][0])[:32]:
    print('| %.1f%% | %.2f | %d | `%s` |' % (100 * ms / total, ms, c, name))
ours = sum(ms for k, (ms, c) in rows.items() if not (k.startswith('at::') or 'cublas' in k or 'cutlass' in k or 'gemv' in k
                                                      or k.startswith('std::') or 'internal::' in k or 'dot_kernel' in k
                                                      or 'reduce_1Block' in k or 'gemmSN' in k))
conv = sum(ms for k, (ms, c) in rows.items() if k.startswith('conv_igemm_kernel') or k.startswith('conv_wgrad_kernel'))
print('\nTotal kernel time over %d launches: %.1f ms. Kernels of libimx_kernels.so: %.1f ms (%.1f%%), of which the tcgen05 '
      'implicit-GEMM family (conv_igemm / conv_wgrad): %.1f ms (%.1f%%).' % (n, total, ours, 100 * ours / total, conv, 100 * conv / total))
