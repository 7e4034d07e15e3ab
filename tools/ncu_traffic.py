"""Writes profiles/r02_conv_traffic.json from an `ncu --set full` capture of the dominant convolution launch
(conv_igemm_kernel<256,0,1>, 3x3 512 -> 512 at 8 x 128 x 256: `python tools/ncu_kernels.py conv`).

    ncu --set full --clock-control none --import-source on -k regex:conv_igemm_kernel -s 1 -c 1 \
        -o gpurun_out/prof_conv_r02 -f python tools/ncu_kernels.py conv          (on the GPU box)
    python tools/ncu_traffic.py gpurun_out/prof_conv_r02.ncu-rep                  (here)

bench.py reads the JSON for roofline.traffic (bytes per launch)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, body = rows[0], rows[1], rows[2:]
    i = {n: k for k, n in enumerate(head)}
    r = body[0]

    def val(name):
        v, u = float(r[i[name]]), units[i[name]]
        return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}.get(u, 1.0)
    d = {'kernel': r[i['Kernel Name']], 'source': os.path.basename(path),
         'dram_bytes_read': val('dram__bytes_read.sum'), 'dram_bytes_write': val('dram__bytes_write.sum'),
         'duration_us': float(r[i['gpu__time_duration.sum']]) * {'msecond': 1e3, 'usecond': 1.0, 'nsecond': 1e-3}.get(units[i['gpu__time_duration.sum']], 1.0),
         'tensor_pipe_active_pct': float(r[i['sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active']])
         if 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active' in i else None,
         'algorithmic_bytes': 8 * 128 * 256 * 512 * 2 * 2 + 512 * 512 * 9 * 2,
         'workload': '3x3 conv 512 -> 512 on 8 x 128 x 256 bf16 NHWC (tools/ncu_kernels.py conv), one launch, ncu --set full'}
    with open(os.path.join(ROOT, 'profiles', 'r02_conv_traffic.json'), 'w') as f:
        json.dump(d, f, indent=1)
    print(json.dumps(d, indent=1))


if __name__ == '__main__':
    main(sys.argv[1])
