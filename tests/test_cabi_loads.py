"""The C-ABI library builds for sm_100a, loads, and exports every symbol that
include/mlc.h declares (no compute here: no GPU in the build container)."""
import os
import re

import pytest


def test_library_exports_every_declared_symbol():
    from mlcomp_b200 import _lib
    lib = _lib.load()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, 'include', 'mlc.h')).read()
    declared = set(re.findall(r'\b(mlc_[a-z0-9_]+)\s*\(', header))
    assert declared, 'no declarations found in mlc.h'
    assert declared == set(_lib.SIGNATURES), \
        'ctypes table and mlc.h disagree: %s' % (declared ^ set(_lib.SIGNATURES))
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.mlc_version() >= 100
    assert lib.mlc_last_error() is not None


def test_host_side_argument_errors_need_no_gpu():
    """Argument validation happens before any CUDA call."""
    import ctypes
    from mlcomp_b200 import _lib
    lib = _lib.load()
    thr = (ctypes.c_float * 2)(0.5, 0.25)
    rc = lib.mlc_seg_lut_build(thr, 2, _lib.MLC_ACT_SIGMOID,
                               ctypes.c_void_p(16), None)
    assert rc == _lib.MLC_ERR_ARG
    assert b'ascending' in lib.mlc_last_error()
    rc = lib.mlc_confusion(None, 0, 5, 3, None, None, None, None, None)
    assert rc == _lib.MLC_ERR_ARG
    assert lib.mlc_dice_search_workspace(6666, 4, 100, 5) == \
        148 * 4 * 100 * 5 * 8


def test_ops_refuse_cpu_tensors():
    import torch
    from mlcomp_b200 import ops
    with pytest.raises(TypeError, match='CUDA tensor'):
        ops.confusion(torch.zeros(4, 3), torch.zeros(4, dtype=torch.int64))


def test_compose_tta_is_a_dihedral_element():
    import itertools
    import torch
    import oracle
    from mlcomp_b200 import ops
    from mlcomp_b200._lib import MLC_XF_HFLIP, MLC_XF_TRANSPOSE, MLC_XF_VFLIP
    h, w = 3, 5
    for k in range(4):
        for tags in itertools.product(oracle.TTA_TAGS, repeat=k):
            xf = ops.compose_tta(tags)
            tr = bool(xf & MLC_XF_TRANSPOSE)
            hs, ws = (w, h) if tr else (h, w)
            src = torch.arange(hs * ws).reshape(1, 1, hs, ws)
            if oracle.tta_inverse(src, tags).shape[2:] != (h, w):
                continue
            want = oracle.tta_inverse(src, tags)[0, 0]
            got = torch.empty(h, w, dtype=src.dtype)
            for y in range(h):
                for x in range(w):
                    u, v = (x, y) if tr else (y, x)
                    ys = hs - 1 - u if xf & MLC_XF_VFLIP else u
                    xs = ws - 1 - v if xf & MLC_XF_HFLIP else v
                    got[y, x] = src[0, 0, ys, xs]
            assert torch.equal(got, want), tags
