"""Builds oracle/_build/libresample_oracle.so from oracle/resample_oracle.c with gcc.

TEST INFRASTRUCTURE ONLY.  Called by __graft_entry__.build() and lazily by
oracle/warp_oracle.py.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'resample_oracle.c')
OUT_DIR = os.path.join(HERE, '_build')
LIB = os.path.join(OUT_DIR, 'libresample_oracle.so')


def build(force=False):
    os.makedirs(OUT_DIR, exist_ok=True)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        cmd = ['gcc', '-O2', '-fPIC', '-shared', '-ffp-contract=off', '-fopenmp', SRC, '-o', LIB, '-lm']
        subprocess.run(cmd, check=True)
    return LIB


if __name__ == '__main__':
    print(build(force=True))
