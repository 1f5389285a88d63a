"""Tensors that begin / end exactly at the edges of their allocation, and at
odd element offsets inside it: the window loads of the segmentation kernels
read whole aligned words around an item and must never leave the tensor.  On
the default library these are parity tests; run with the bounds-checked build
(``python -m mlcomp_b200.build --bounds-check``, then
``MLC_LIB=mlcomp_b200/lib/libmlcomp_b200_chk.so pytest -m gpu``) every load of
every GPU test is checked against the registered tensor ranges and a stray one
traps (compute-sanitizer is closed on this pool)."""
import numpy as np
import pytest
import torch

import oracle
from test_gpu_parity import DEV, forward_tta, synth_seg

pytestmark = pytest.mark.gpu
FLIPS = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]


@pytest.fixture(scope='module')
def ops():
    from mlcomp_b200 import ops as o
    return o


def at_edge(t, where):
    """A copy of ``t`` that is a view into a fresh allocation: 'start' = the
    allocation begins with it, 'end' = the allocation ends with it, 'odd' =
    one element into the allocation (2-byte aligned for fp16)."""
    n = t.numel()
    pad = 0 if where == 'start' else (1 if where == 'odd' else 7)
    buf = torch.empty(n + pad, dtype=t.dtype, device=DEV)
    view = buf[pad:] if where != 'start' else buf[:n]
    view.copy_(t.reshape(-1).to(DEV))
    return view.view(t.shape)


@pytest.mark.parametrize('where', ['start', 'end', 'odd'])
@pytest.mark.parametrize('dtype', [torch.float16, torch.float32])
@pytest.mark.parametrize('shape', [(2, 2, 35, 53), (1, 3, 16, 40),
                                   (2, 1, 7, 9), (1, 1, 64, 96)])
def test_hist_prob_rle_at_allocation_edges(ops, where, dtype, shape):
    n, c, h, w = shape
    srcs, gt = synth_seg(n, c, h, w, seed=h * w, dtype=dtype, n_src=4)
    seen = [at_edge(forward_tta(s, t), where) for s, t in zip(srcs, FLIPS)]
    g = at_edge(gt, where)
    base = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, FLIPS)]
    thr = oracle.threshold_grid(100)
    want_p = ops.seg_prob(base, xforms=FLIPS)
    assert torch.equal(ops.seg_prob(seen, xforms=FLIPS), want_p)
    want_h = ops.seg_hist(base, gt.to(DEV), thr, xforms=FLIPS)
    assert torch.equal(ops.seg_hist(seen, g, thr, xforms=FLIPS), want_h)
    p = want_p.cpu().numpy()
    for i in range(n):
        for j in range(c):
            assert np.array_equal(
                want_h[i, j].cpu().numpy(),
                oracle.bin_counts(gt[i, j].numpy(), p[i, j], thr))
    a = ops.rle_encode(seen, 0.4, 3, xforms=FLIPS)
    b = ops.rle_encode(base, 0.4, 3, xforms=FLIPS)
    assert ops.rle_strings(*a[:3]) == ops.rle_strings(*b[:3])
    # single source: table / logit-domain / code paths
    one = ops.seg_hist([seen[0]], g, thr, lut=ops.seg_lut(thr)
                       if dtype == torch.float16 else None)
    assert torch.equal(one, ops.seg_hist([base[0]], gt.to(DEV), thr))
    a1 = ops.rle_encode([seen[0]], 0.5, 0)
    b1 = ops.rle_encode([base[0]], 0.5, 0)
    assert ops.rle_strings(*a1[:3]) == ops.rle_strings(*b1[:3])
