"""GPU parity of seg_hist_guarded (csrc/seg_hist_guarded.cu): the multi-source
fp16 sigmoid histogram that decides most pixels from a bounded-error estimate
and the rest exactly.  Its results must equal, bit for bit, the binning of the
exactly fused probabilities (mlc_seg_prob, itself bit-equal to torch's CUDA
sigmoid + the oracle's fp32 mean): histograms AND per-pixel bin planes.
Reference lines behind the arithmetic: utils/torch.py:15,24-29 (sigmoid on the
device, then TtaWrap.inverse), contrib/transform/tta.py:20-31,
contrib/metrics/dice.py:6-11 (`>=`, the sums)."""
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


def device_bins(ops, srcs, tags, thr):
    """cnt(p) of the exactly fused probabilities, on the device:
    #{k : p >= thr[k]} = searchsorted(thr, p, right=True); NaN -> 0."""
    p = ops.seg_prob(srcs, xforms=tags)
    t = torch.from_numpy(np.asarray(thr, np.float32)).to(p.device)
    c = torch.searchsorted(t, p.reshape(-1), right=True).reshape(p.shape)
    c[torch.isnan(p)] = 0
    return c


def check(ops, srcs, tags, gt, thr):
    hist, bins = ops.seg_hist(srcs, gt, thr, xforms=tags, bins=True)
    hist2 = ops.seg_hist(srcs, gt, thr, xforms=tags)
    assert torch.equal(hist, hist2)
    want = device_bins(ops, srcs, tags, thr)
    assert torch.equal(bins.long(), want), \
        int((bins.long() != want).sum())
    nb = len(thr) + 1
    n, c = gt.shape[:2]
    flat = want.reshape(n * c, -1) + nb * (gt.reshape(n * c, -1) != 0).long()
    ref = torch.zeros((n * c, 2 * nb), dtype=torch.long, device=gt.device)
    ref.scatter_add_(1, flat, torch.ones_like(flat))
    assert torch.equal(hist.reshape(n * c, 2 * nb).long(), ref)
    return hist


@pytest.mark.parametrize('n_src', [2, 3, 4, 5, 8, 9])
@pytest.mark.parametrize('flips', [False, True])
def test_guarded_matches_exact_binning(ops, n_src, flips):
    srcs, gt = synth_seg(3, 2, 64, 96, seed=300 + n_src, n_src=n_src)
    tags = [FLIPS[k % 4] if flips else [] for k in range(n_src)]
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, tags)]
    thr = oracle.threshold_grid(100)
    hist = check(ops, seen, tags, gt.to(DEV), thr)
    # and against the oracle's own binning of those probabilities on the CPU
    p = ops.seg_prob(seen, xforms=tags).cpu().numpy()
    for i in range(gt.shape[0]):
        for j in range(gt.shape[1]):
            assert np.array_equal(
                hist[i, j].cpu().numpy(),
                oracle.bin_counts(gt[i, j].numpy(), p[i, j], thr))


def test_guarded_every_fp16_code(ops):
    """All 65 536 codes (NaN, +-inf, +-65504, denormals, +-0) as S equal
    sources and against shifted copies: overflow of the denominator product
    and NaN items must take the exact path."""
    codes = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    x = codes.view(torch.float16).reshape(1, 1, 256, 256).to(DEV)
    gt = (torch.arange(65536).reshape(1, 1, 256, 256) % 3 == 0) \
        .to(torch.uint8).to(DEV)
    thr = oracle.threshold_grid(100)
    for n_src in (2, 4, 5):
        check(ops, [x] * n_src, [[]] * n_src, gt, thr)
    rolled = [x, x.flatten().roll(4099).reshape(x.shape),
              x.flatten().roll(17).reshape(x.shape),
              x.flatten().roll(33333).reshape(x.shape)]
    rolled = [r.contiguous() for r in rolled]
    check(ops, rolled, [[]] * 4, gt, thr)
    check(ops, rolled, FLIPS, gt, thr)


def test_guarded_means_on_and_next_to_thresholds(ops):
    """Sources whose exact mean lands on a threshold or 1 ulp beside it: the
    estimate cannot decide those pixels, the exact path must."""
    thr = oracle.threshold_grid(100)
    halves = torch.arange(-0x4C00, 0x4C00, dtype=torch.int32)
    vals = torch.where(halves < 0, (-halves) | 0x8000, halves).to(torch.int16) \
        .view(torch.float16)                      # every |x| < 16
    sig = torch.sigmoid(vals.float().to(DEV))     # the device's exact sigmoid
    t = torch.from_numpy(thr).to(DEV)
    # for every threshold: pairs (x, x') with (s(x) + s(x')) / 2 nearest to it
    rows = []
    for k in range(1, 100):
        target = 2 * t[k] - sig[::37]             # wanted s(x')
        j = torch.searchsorted(sig, target).clamp(1, len(sig) - 2)
        for d in (-1, 0, 1):
            rows.append(torch.stack([vals[::37].to(DEV), vals.to(DEV)[j + d]]))
    pairs = torch.cat(rows, dim=1)                # [2, n]
    n = pairs.shape[1] // 8 * 8
    a = pairs[0, :n].reshape(1, 1, 1, n).contiguous()
    b = pairs[1, :n].reshape(1, 1, 1, n).contiguous()
    gt = torch.zeros((1, 1, 1, n), dtype=torch.uint8, device=DEV)
    check(ops, [a, b], [[], []], gt, thr)
    check(ops, [a, b, b, a], [[], []] * 2, gt, thr)


@pytest.mark.parametrize('thr', [
    np.float32([0.5]),                                  # one threshold
    np.float32([0.1, 0.11, 0.5, 0.9]),                  # irregular
    np.float32(np.arange(50) / 100 + 0.25),             # does not span [0, 1]
    np.float32(np.arange(200) / 200),                   # more than 120
    np.float32(np.arange(10) / 10),                     # coarse uniform
], ids=['one', 'irregular', 'partial', 'fine', 'coarse'])
def test_guarded_or_fallback_any_table(ops, thr):
    """Grids the guarded kernel declines stay on the exact kernels; either
    way the result is the exact binning."""
    srcs, gt = synth_seg(2, 2, 32, 64, seed=41, n_src=4)
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, FLIPS)]
    check(ops, seen, FLIPS, gt.to(DEV), thr)


def test_guarded_severstal_tta_full_planes(ops):
    """4-flip TTA on 256x1600 planes (26 M pixels): every pixel's bin equals
    the binning of the exact probabilities; the dice table equals the one
    computed from the exact histograms."""
    g = torch.Generator(device=DEV).manual_seed(5)
    n, c, h, w = 16, 4, 256, 1600
    base = torch.randn((n, c, h, w), generator=g, device=DEV) * 1.5 - 4
    base[:, :, 64:128, 200:900] += 8
    gt = torch.zeros((n, c, h, w), dtype=torch.uint8, device=DEV)
    gt[:, :, 64:128, 200:900] = 1
    gt[::2] = 0
    srcs = []
    for dims in ((), (3,), (2,), (2, 3)):
        v = (base + 0.5 * torch.randn(base.shape, generator=g, device=DEV)) \
            .half()
        srcs.append(v.flip(*dims).contiguous() if dims else v)
    thr = oracle.threshold_grid(100)
    check(ops, srcs, FLIPS, gt, thr)
    # a sharper set: confident logits N(+-6, 1)
    sharp = [(s.float() * 0.66 + (s.float() > 0) * 4 - 2).half() for s in srcs]
    check(ops, sharp, FLIPS, gt, thr)


def test_guarded_split_planes_small_job(ops):
    """Few large planes: the host splits planes into units (atomic flush)."""
    g = torch.Generator(device=DEV).manual_seed(9)
    x = [(torch.randn((1, 1, 1024, 2048), generator=g, device=DEV) * 3).half()
         for _ in range(4)]
    gt = (torch.rand((1, 1, 1024, 2048), generator=g, device=DEV) < 0.3) \
        .to(torch.uint8)
    check(ops, x, [[]] * 4, gt, oracle.threshold_grid(100))
    check(ops, x, FLIPS, gt, oracle.threshold_grid(100))
