"""Build libdca_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB_PATH = os.path.join(HERE, 'libdca_b200.so')
SOURCES = ['dca_b200.cu']
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith('.cuh')) + [os.path.join('..', '..', 'include', 'dca_b200.h')]
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              # numpy evaluates every product and sum with its own rounding: never contract into FMA
              '-fmad=false', '-shared', '-Xcompiler', '-fPIC']


def _nvcc():
    for p in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
        if p and (os.path.sep not in p or os.path.exists(p)):
            return p
    return 'nvcc'


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + \
          ['-o', LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd)
    return LIB_PATH


if __name__ == '__main__':
    import sys
    print(build(force=True, verbose="-v" in sys.argv))
