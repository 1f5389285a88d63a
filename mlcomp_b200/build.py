"""Build libmlcomp_b200.so in-tree with nvcc for sm_100a (and nothing else).

    python -m mlcomp_b200.build [--verbose] [--dev-knobs] [--bounds-check]
    python -m mlcomp_b200.build --variant NAME --define A=B [--define C ...]

The library is a plain C-ABI shared object (include/mlc.h): no torch headers,
no pybind.  It is git-ignored but travels to the GPU box with the snapshot.

Every .cu of csrc/ is compiled to its own object (in parallel, rebuilt only
when it or a header is newer) and the objects are linked into one library.
Variants (separate object directories and library names, so they can sit
side by side):
  --dev-knobs     -DMLC_DEV_KNOBS: the retired kernel schedules and the
                  MLC_* environment switches used for A/B measurements
  --bounds-check  -DMLC_BOUNDS_CHECK: every global load of the segmentation
                  kernels is checked against [tensor, tensor + bytes) and
                  traps when outside (libmlcomp_b200_chk.so; run the GPU
                  suite with MLC_LIB=.../libmlcomp_b200_chk.so)
  --variant NAME  any other set of --define's -> libmlcomp_b200_NAME.so (A/B
                  builds of compile-time tuning constants)
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB_DIR = os.path.join(HERE, 'lib')
LIB = os.path.join(LIB_DIR, 'libmlcomp_b200.so')
SOURCES = ['common.cu', 'confusion.cu', 'seg_hist.cu', 'seg_hist_guarded.cu',
           'seg_hist_guarded_rows.cu', 'seg_argmax.cu', 'dice_scan.cu', 'rle.cu', 'rle_format.cu', 'rle_decode.cu',
           'composite.cu', 'exchange.cu']


def nvcc_path():
    for cand in (os.environ.get('NVCC'), shutil.which('nvcc'),
                 '/usr/local/cuda/bin/nvcc'):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError('nvcc not found (set NVCC=/path/to/nvcc)')


def lib_for(variant=''):
    return os.path.join(LIB_DIR, 'libmlcomp_b200%s.so' %
                        ('_' + variant if variant else ''))


def _headers():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)
            if f.endswith(('.cuh', '.h'))]
    deps.append(os.path.join(HERE, '..', 'include', 'mlc.h'))
    return deps


def _digest(flag_text=''):
    """Content hash of everything the library is built from (not mtimes: a
    snapshot copied to another machine must not look stale)."""
    import hashlib
    h = hashlib.sha256(flag_text.encode())
    for d in sorted([os.path.join(CSRC, s) for s in SOURCES] + _headers()):
        h.update(os.path.basename(d).encode())
        with open(d, 'rb') as f:
            h.update(f.read())
    return h.hexdigest()


def _base_flags(defines=()):
    return ' '.join(['-O3', '-std=c++17', '-lineinfo', '-gencode',
                     'arch=compute_100a,code=sm_100a'] +
                    ['-D' + d for d in defines] +
                    os.environ.get('MLC_EXTRA_NVCC', '').split())


def needs_build(variant='', defines=()):
    lib = lib_for(variant)
    stamp = lib + '.stamp'
    if not os.path.exists(lib) or not os.path.exists(stamp):
        return True
    return open(stamp).read().strip() != _digest(_base_flags(defines))


def build(verbose=False, force=False, variant='', defines=()):
    """Compile what is stale and link.  Returns the library path."""
    lib = lib_for(variant)
    if not force and not needs_build(variant, defines):
        return lib
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(HERE, 'build', variant or 'default')
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = nvcc_path()
    flags = ['-O3', '-std=c++17', '-lineinfo',
             '-gencode', 'arch=compute_100a,code=sm_100a',
             '-Xcompiler', '-fPIC,-O3,-Wall']
    flags += ['-D' + d for d in defines]
    if verbose:
        flags += ['-Xptxas', '-v']
    # A/B builds of compile-time knobs (e.g. MLC_EXTRA_NVCC='-DMLC_ROW_BATCH=2')
    extra = os.environ.get('MLC_EXTRA_NVCC', '').split()
    flags += extra
    stamp = os.path.join(obj_dir, 'flags.txt')
    flag_text = ' '.join(flags)
    if not os.path.exists(stamp) or open(stamp).read() != flag_text:
        force = True
    hdr_t = max(os.path.getmtime(h) for h in _headers())

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(obj_dir, src[:-3] + '.o')
        if not force and os.path.exists(o) and \
                os.path.getmtime(o) > max(os.path.getmtime(s), hdr_t):
            return o, '', 0
        res = subprocess.run([nvcc, '-c', s, '-o', o] + flags,
                             capture_output=True, text=True)
        return o, res.stdout + res.stderr, res.returncode

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_one, SOURCES))
    failed = False
    for o, log, rc in results:
        if log and (verbose or rc):
            sys.stderr.write(log)
        failed |= rc != 0
    if failed:
        raise RuntimeError('nvcc failed building libmlcomp_b200.so')
    open(stamp, 'w').write(flag_text)
    res = subprocess.run([nvcc, '-shared', '-o', lib, '-cudart', 'static',
                          '-gencode', 'arch=compute_100a,code=sm_100a'] +
                         [o for o, _, _ in results],
                         capture_output=True, text=True)
    if verbose or res.returncode:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode:
        raise RuntimeError('nvcc failed linking libmlcomp_b200.so')
    with open(lib + '.stamp', 'w') as f:
        f.write(_digest(_base_flags(defines)))
    return lib


def build_variant(name, verbose=False, force=False):
    defs = {'dev': ('MLC_DEV_KNOBS',), 'chk': ('MLC_BOUNDS_CHECK',)}[name]
    return build(verbose=verbose, force=force, variant=name, defines=defs)


if __name__ == '__main__':
    v = '--verbose' in sys.argv
    if '--variant' in sys.argv:
        name = sys.argv[sys.argv.index('--variant') + 1]
        defs = [sys.argv[i + 1] for i, x in enumerate(sys.argv)
                if x == '--define']
        print(build(verbose=v, force=True, variant=name, defines=defs))
    elif '--dev-knobs' in sys.argv:
        print(build_variant('dev', verbose=v, force=True))
    elif '--bounds-check' in sys.argv:
        print(build_variant('chk', verbose=v, force=True))
    else:
        print(build(verbose=v, force=True))
