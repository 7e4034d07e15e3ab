"""Builds oracle/_ref/libref_cuda_kernels.so: the REFERENCE's own CUDA kernel bodies, compiled for host cores.

TEST INFRASTRUCTURE ONLY.  Works only where /root/reference exists (the build container); the GPU box
uses the prebuilt library that travels with the snapshot, and the committed fixtures under tests/golden/.

Recipe (nothing from the reference is copied into the repository; intermediates and the library go to
oracle/_ref/, which is git-ignored):
  1. read each kernel source where it lies under /root/reference,
  2. keep only the device code: everything before the first host launcher (those use the `<<< >>>` launch
     syntax and the pre-1.5 at::Tensor API, which are not C++ / do not exist any more),
  3. `static __shared__` -> `__shared__` (the shim maps `__shared__` to static storage already),
  4. compile the result with g++ against oracle/ref_cuda/shim/cuda_on_host.h together with the host
     launchers in this directory, twice:
       libref_cuda_kernels.so       -ffp-contract=off: the source's arithmetic as C++ defines it, every
                                    product and sum rounded separately;
       libref_cuda_kernels_fmad.so  -ffp-contract=fast -march=haswell: `acc += a * b` becomes one fused multiply-add,
                                    which is what nvcc's default -fmad=true does to the same source (checked
                                    against the PTX nvcc 12.9 emits for these kernels: the all-float products
                                    feeding an accumulation are contracted, e.g. the fourth tap of the resample2d
                                    forward and every term of its flow gradient; products with a double operand
                                    are not).
     The two differ by at most one rounding per accumulated term; tests state which variant a kernel equals bit
     for bit and hold it to the other within a tight tolerance.

    python -m oracle.ref_cuda.build_ref
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE = os.path.dirname(HERE)
OUT = os.path.join(ORACLE, '_ref')
GEN = os.path.join(OUT, 'gen')
LIB = os.path.join(OUT, 'libref_cuda_kernels.so')
LIB_FMAD = os.path.join(OUT, 'libref_cuda_kernels_fmad.so')
REF_TP = '/root/reference/imaginaire/third_party'

# (source under the reference tree, first line of its first host launcher, host-launcher file here)
UNITS = [
    ('resample2d/src/resample2d_kernel.cu', 'void resample2d_kernel_forward(', 'ref_resample2d_host.cpp'),
    ('channelnorm/src/channelnorm_kernel.cu', 'void channelnorm_kernel_forward(', 'ref_channelnorm_host.cpp'),
    ('correlation/src/correlation_cuda_kernel.cu', 'int correlation_forward_cuda_kernel(', 'ref_correlation_host.cpp'),
]


def available():
    return os.path.isdir(REF_TP)


def _device_part(path, cut_marker):
    lines = open(path).read().split('\n')
    for i, line in enumerate(lines):
        if line.startswith(cut_marker):
            break
    else:
        raise RuntimeError('%s: host launcher %r not found' % (path, cut_marker))
    return '\n'.join(lines[:i]).replace('static __shared__', '__shared__') + '\n'


def build(force=False):
    """Returns the path of the -ffp-contract=off library (its _fmad sibling lies beside it), or None when neither the
    reference nor a prebuilt library is present."""
    if not available():
        return LIB if os.path.exists(LIB) else None
    os.makedirs(GEN, exist_ok=True)
    srcs = [os.path.join(HERE, u[2]) for u in UNITS] + [os.path.join(HERE, 'shim', 'cuda_on_host.h'), __file__]
    srcs += [os.path.join(REF_TP, u[0]) for u in UNITS]
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(s) for s in srcs):
        return LIB
    for rel, marker, host in UNITS:
        inc = os.path.join(GEN, os.path.basename(rel)[:-3] + '.device.inc')
        with open(inc, 'w') as f:
            f.write(_device_part(os.path.join(REF_TP, rel), marker))
    for lib, tag, fp_flags in ((LIB_FMAD, 'fmad', ['-ffp-contract=fast', '-march=haswell']), (LIB, 'strict', ['-ffp-contract=off'])):
        objs = []
        for rel, marker, host in UNITS:
            obj = os.path.join(GEN, '%s.%s.o' % (host[:-4], tag))
            cmd = ['g++', '-std=c++17', '-O2', '-fPIC', '-Wno-unknown-pragmas', '-Wno-unused-variable'] + fp_flags + \
                  ['-I', os.path.join(HERE, 'shim'), '-I', GEN, '-I', os.path.dirname(os.path.join(REF_TP, rel)),
                   '-c', os.path.join(HERE, host), '-o', obj]
            subprocess.run(cmd, check=True)
            objs.append(obj)
        subprocess.run(['g++', '-shared', '-o', lib] + objs, check=True)
    return LIB


if __name__ == '__main__':
    print(build(force=True))
