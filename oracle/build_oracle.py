"""Builds oracle/_build/libresample_oracle.so and libflownet_ops_oracle.so from the C restatements with gcc.

TEST INFRASTRUCTURE ONLY.  Called by __graft_entry__.build() and lazily by
oracle/warp_oracle.py.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'resample_oracle.c')
OUT_DIR = os.path.join(HERE, '_build')
LIB = os.path.join(OUT_DIR, 'libresample_oracle.so')
SRC_FLOWNET = os.path.join(HERE, 'flownet_ops_oracle.c')
LIB_FLOWNET = os.path.join(OUT_DIR, 'libflownet_ops_oracle.so')


def _compile(src, lib, force):
    os.makedirs(OUT_DIR, exist_ok=True)
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(src):
        cmd = ['gcc', '-O2', '-fPIC', '-shared', '-ffp-contract=off', '-fopenmp', src, '-o', lib, '-lm']
        subprocess.run(cmd, check=True)
    return lib


def build(force=False):
    _compile(SRC_FLOWNET, LIB_FLOWNET, force)
    return _compile(SRC, LIB, force)


def build_flownet(force=False):
    return _compile(SRC_FLOWNET, LIB_FLOWNET, force)


if __name__ == '__main__':
    print(build(force=True))
