"""Summarises an .ncu-rep (ncu --set full) into a markdown table for profiles/.

Usage: python tools/ncu_summarise.py gpurun_out/prof_conv_r01.ncu-rep > profiles/r01_ncu_conv_summary.md
"""
import csv
import io
import subprocess
import sys

METRICS = [
    ('gpu__time_duration.sum', 'dur'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm%'),
    ('sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'tensor%'),
    ('sm__inst_executed_pipe_tensor.sum', 'tensor_inst'),
    ('dram__throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
    ('dram__bytes_read.sum', 'dram_rd'),
    ('dram__bytes_write.sum', 'dram_wr'),
    ('lts__t_sector_hit_rate.pct', 'L2hit%'),
    ('l1tex__throughput.avg.pct_of_peak_sustained_active', 'L1%'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'),
    ('launch__registers_per_thread', 'regs'),
    ('launch__grid_size', 'grid'),
    ('launch__block_size', 'block'),
]


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, body = rows[0], rows[1], rows[2:]
    idx = {n: i for i, n in enumerate(head)}
    cols = [(m, s) for m, s in METRICS if m in idx]
    extra = [n for n in head if 'tensor' in n and 'pct' in n]
    print('# ncu --set full summary of `%s`\n' % path.split('/')[-1])
    print('Columns: ' + ', '.join('%s=%s [%s]' % (s, m, units[idx[m]]) for m, s in cols) + '\n')
    print('| # | kernel | ' + ' | '.join(s for _, s in cols) + ' |')
    print('|---|---|' + '---|' * len(cols))
    for r in body:
        name = r[idx['Kernel Name']]
        name = name.replace('imx::', '')[:70]
        print('| %s | `%s` | ' % (r[idx['ID']], name) + ' | '.join(r[idx[m]] for m, _ in cols) + ' |')
    if extra:
        print('\nTensor-pipe metrics present in the report: ' + ', '.join(extra[:12]))


if __name__ == '__main__':
    main(sys.argv[1])
