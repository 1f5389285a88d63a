"""mlc_dice_search_static: the dice search for callers that cannot allocate.

Some users of libmlcomp_b200 have input and output buffers but no way to get a
workspace: a scoring loop that shares the card with an inference engine which
reserved all device memory at start-up, or a ctypes binding from the
reference's worker process that owns no device allocator.  For them the
library carries a fixed 4 MiB scratch array; this file checks that the entry
point using it returns exactly what mlc_dice_search returns with a
caller-allocated workspace - twice in a row with different inputs, because the
array is shared by all calls and keeps the previous call's contents - and that
a search too large for the array is refused with a clear message.

Named test_zz_* so that pytest collects it last.
"""
import ctypes
from ctypes import POINTER, c_int32, c_void_p

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = 'cuda:0'


def _call_static(lib, hist, areas, empty_one=True, eps=1e-6):
    """Straight through the C ABI: torch only owns the in/out buffers."""
    n, c, _, nb = hist.shape
    b = nb - 1
    areas = np.ascontiguousarray(areas, dtype=np.int32)
    a = areas.shape[1]
    out = torch.full((c, b, a), float('nan'), dtype=torch.float64, device=DEV)
    micro = torch.full((c, 1 + 2 * b), -1, dtype=torch.int64, device=DEV)
    rc = lib.mlc_dice_search_static(
        c_void_p(hist.data_ptr()), n, c, b,
        areas.ctypes.data_as(POINTER(c_int32)), a, int(empty_one), eps,
        c_void_p(out.data_ptr()), c_void_p(micro.data_ptr()),
        c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, out, micro


def _hist(seed, n, c, b):
    rng = np.random.default_rng(seed)
    h = rng.integers(0, 4000, (n, c, 2, b + 1)).astype(np.int32)
    h[::3, :, 1, :] = 0                 # planes without ground truth
    h[::5, :, :, 1:] = 0                # planes with nothing above thr[0]
    return torch.from_numpy(h).to(DEV)


def test_static_scratch_matches_workspace_version_twice():
    from mlcomp_b200 import _lib, ops
    lib = _lib.load()
    torch.cuda.set_device(0)
    # call 1: the C2 grid; call 2: other shape and values, so anything left
    # in the shared array by call 1 would show up
    cases = [(1, 900, 4, 100, [[0, 256, 512, 1024, 2048]] * 4),
             (2, 37, 3, 64, [[0, 10], [5, 500], [0, 10 ** 6]])]
    for seed, n, c, b, areas in cases:
        hist = _hist(seed, n, c, b)
        want, want_micro = ops.dice_search(hist, areas)
        rc, got, micro = _call_static(lib, hist, areas)
        assert rc == 0, lib.mlc_last_error()
        torch.cuda.synchronize()
        # same kernels, same summation order: equal to the last bit (the
        # other parity tests allow 1e-5 relative; this one needs none)
        assert torch.equal(got, want)
        assert torch.equal(micro, want_micro)
    # and once more with the first shape: stale partials of the smaller
    # second search must not leak into a larger one
    seed, n, c, b, areas = cases[0]
    hist = _hist(7, n, c, b)
    want, _ = ops.dice_search(hist, areas, empty_one=False, eps=1e-3)
    rc, got, _ = _call_static(lib, hist, areas, empty_one=False, eps=1e-3)
    assert rc == 0, lib.mlc_last_error()
    assert torch.equal(got, want)


def test_static_scratch_refuses_searches_beyond_its_size():
    from mlcomp_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    n, c, b = 2000, 8, 255
    areas = [[0, 1, 2, 3, 4, 5, 6, 7]] * c          # 8*255*8 = 16 320 cells
    need = lib.mlc_dice_search_workspace(n, c, b, 8)
    assert need > 4 * 1024 * 1024
    hist = torch.zeros((n, c, 2, b + 1), dtype=torch.int32, device=DEV)
    rc, out, _ = _call_static(lib, hist, areas)
    assert rc == _lib.MLC_ERR_CAPACITY
    msg = lib.mlc_last_error().decode()
    assert 'built-in array' in msg and str(need) in msg
    torch.cuda.synchronize()
    assert torch.isnan(out).all()                   # nothing was launched
