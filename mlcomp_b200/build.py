"""Build libmlcomp_b200.so in-tree with nvcc for sm_100a (and nothing else).

    python -m mlcomp_b200.build [--verbose]

The library is a plain C-ABI shared object (include/mlc.h): no torch headers,
no pybind.  It is git-ignored but travels to the GPU box with the snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB_DIR = os.path.join(HERE, 'lib')
LIB = os.path.join(LIB_DIR, 'libmlcomp_b200.so')
SOURCES = ['common.cu', 'confusion.cu', 'seg_hist.cu', 'dice_scan.cu',
           'rle.cu', 'rle_format.cu', 'rle_decode.cu', 'composite.cu']


def nvcc_path():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'),
                 '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found (set NVCC=/path/to/nvcc)')


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, '..', 'include', 'mlc.h'))
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose=False, force=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    cmd = [nvcc_path(), '-shared', '-o', LIB, '-O3', '-std=c++17', '-lineinfo',
           '-gencode', 'arch=compute_100a,code=sm_100a',
           '-Xcompiler', '-fPIC,-O3,-Wall', '-cudart', 'static']
    if verbose:
        cmd += ['-Xptxas', '-v']
    # A/B builds of compile-time knobs (e.g. MLC_EXTRA_NVCC='-DMLC_ROW_BATCH=2')
    cmd += os.environ.get('MLC_EXTRA_NVCC', '').split()
    cmd += [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode:
        raise RuntimeError('nvcc failed building libmlcomp_b200.so')
    return LIB


if __name__ == '__main__':
    print(build(verbose='--verbose' in sys.argv, force=True))
