"""GPU: full plane shapes of the BASELINE configs, checked through properties
that need no CPU pass over the pixels (the oracle would take core-hours at
these sizes): conservation of pixel counts, monotone cumulative counts,
encode -> decode round trips, shard-sum == whole, and spot planes against the
oracle."""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _synth(n, c, h, w, seed, n_src=1, empty=0.5):
    g = torch.Generator(device=DEV).manual_seed(seed)
    yy = torch.arange(h, device=DEV, dtype=torch.float32)[None, :, None]
    xx = torch.arange(w, device=DEV, dtype=torch.float32)[None, None, :]
    p = n * c
    cy = torch.rand(p, generator=g, device=DEV) * h
    cx = torch.rand(p, generator=g, device=DEV) * w
    r = (0.05 + 0.2 * torch.rand(p, generator=g, device=DEV)) * min(h, w)
    keep = torch.rand(p, generator=g, device=DEV) >= empty
    gt = torch.empty((p, h, w), dtype=torch.uint8, device=DEV)
    for lo in range(0, p, 128):
        hi = min(p, lo + 128)
        m = (((yy - cy[lo:hi, None, None]) / r[lo:hi, None, None]) ** 2 +
             ((xx - cx[lo:hi, None, None]) / (2 * r[lo:hi, None, None])) ** 2
             ) <= 1
        gt[lo:hi] = (m & keep[lo:hi, None, None]).to(torch.uint8)
    gt = gt.view(n, c, h, w)
    srcs = []
    for _ in range(n_src):
        x = torch.empty((n, c, h, w), dtype=torch.float16, device=DEV)
        for lo in range(0, n, 32):
            hi = min(n, lo + 32)
            noise = torch.randn((hi - lo, c, h, w), generator=g, device=DEV)
            x[lo:hi] = (gt[lo:hi].float() * 8 - 4 + 1.6 * noise).half()
        srcs.append(x)
    return srcs, gt


def _check_hist_invariants(hist, gt):
    n, c, h, w = gt.shape
    h64 = hist.to(torch.int64)
    assert torch.equal(h64.sum(dim=(2, 3)),
                       torch.full((n, c), h * w, dtype=torch.int64, device=DEV))
    # T from the histogram == ground-truth pixel count (dice.py:10)
    assert torch.equal(h64[:, :, 1].sum(-1),
                       (gt != 0).sum(dim=(2, 3)).to(torch.int64))


def test_c2_shape_table_path_and_sharding():
    """256x1600, 4 classes (C2 / C5 plane shape)."""
    from mlcomp_b200 import ops, scoring
    n, c, h, w = 96, 4, 256, 1600
    (x,), gt = _synth(n, c, h, w, 2, empty=0.86)
    thr = ops.threshold_grid(100)
    areas = [0, 256, 512, 1024, 2048]
    whole = scoring.SegSearch(c, thr, areas)
    hist = whole.add([x], gt)
    _check_hist_invariants(hist, gt)
    # thr[0] = 0: every pixel is predicted at the first threshold
    _, p, i = ops.counts_from_hist(hist)
    assert (p[..., 0] == h * w).all()
    assert (p[..., 1:] <= p[..., :-1]).all() and (i <= p).all()
    # table path == arithmetic path on full-size planes
    assert torch.equal(hist, ops.seg_hist([x], gt, thr))
    # two shards combined == the whole (what the N>1 bench relies on)
    a, b = scoring.SegSearch(c, thr, areas), scoring.SegSearch(c, thr, areas)
    a.add([x[:40]], gt[:40])
    b.add([x[40:].contiguous()], gt[40:].contiguous())
    assert torch.equal(a.micro + b.micro, whole.micro)
    torch.testing.assert_close(a.dice_sum + b.dice_sum, whole.dice_sum,
                               rtol=1e-12, atol=0)
    # spot planes against the CPU oracle
    prob = oracle.fuse_probs([x[:2].cpu()], [[]], 'sigmoid')
    for j in range(c):
        want = oracle.bin_counts(gt[0, j].cpu().numpy(), prob[0, j], thr)
        got = hist[0, j].cpu().numpy()
        assert np.abs(got - want).sum() <= 4          # <= 2 straddling pixels


def test_c3_shape_four_flip_tta():
    """350x525 (rows are not 16-byte multiples), identity / h / v / hv."""
    from mlcomp_b200 import ops
    n, c, h, w = 24, 4, 350, 525
    srcs, gt = _synth(n, c, h, w, 3, n_src=4)
    tags = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]
    seen = [oracle.tta_inverse(s, list(reversed(t))).contiguous()
            for s, t in zip(srcs, tags)]
    thr = ops.threshold_grid(100)
    hist = ops.seg_hist(seen, gt, thr, xforms=tags)
    _check_hist_invariants(hist, gt)
    # flipping every variant back by hand and fusing untransformed sources
    # must give the same histograms
    assert torch.equal(hist, ops.seg_hist(srcs, gt, thr))
    prob = ops.seg_prob(seen, xforms=tags)
    for i, j in ((0, 0), (n - 1, c - 1), (7, 2)):
        want = oracle.bin_counts(gt[i, j].cpu().numpy(),
                                 prob[i, j].cpu().numpy(), thr)
        assert np.array_equal(hist[i, j].cpu().numpy(), want)


def test_c4_shape_rle_round_trip():
    """1024x1024 binary masks: encode -> decode == threshold mask; run lengths
    sum to the area; min-area empties exactly the small masks."""
    from mlcomp_b200 import ops
    from mlcomp_b200.contrib.transform.rle import rle2mask
    n, h, w = 12, 1024, 1024
    (x,), _ = _synth(n, 1, h, w, 4, empty=0.3)
    runs, off, cnt, area = ops.rle_encode([x], 0.5, 3500)
    strings = ops.rle_strings(runs, off, cnt)
    prob = ops.seg_prob([x])
    mask = (prob >= 0.5)
    true_area = mask.sum(dim=(2, 3)).cpu().numpy().ravel()
    assert np.array_equal(area.cpu().numpy().ravel(), true_area)
    for i, s in enumerate(strings):
        if true_area[i] < 3500:
            assert s == ''
            continue
        tok = np.array(s.split(), dtype=np.int64)
        assert tok[1::2].sum() == true_area[i]
        assert (np.diff(tok[0::2]) > 0).all()
        back = rle2mask(s, (w, h))
        assert np.array_equal(back, mask[i, 0].cpu().numpy().astype(np.uint8))
    assert strings[0] == oracle.rle_submission(prob[0, 0].cpu().numpy(), 0.5,
                                               3500)


def test_c5_shape_five_model_ensemble():
    from mlcomp_b200 import ops
    n, c, h, w = 24, 4, 256, 1600
    srcs, gt = _synth(n, c, h, w, 5, n_src=5, empty=0.86)
    thr = ops.threshold_grid(100)
    hist = ops.seg_hist(srcs, gt, thr)
    _check_hist_invariants(hist, gt)
    prob = ops.seg_prob(srcs)
    acc = None
    for s in srcs:
        q = torch.sigmoid(s.float())
        acc = q if acc is None else acc + q
    dev = torch.div(acc, torch.full_like(acc, 5.0))
    assert torch.equal(prob, dev)                 # bit-equal to torch on device
    for i, j in ((0, 0), (n - 1, 3)):
        want = oracle.bin_counts(gt[i, j].cpu().numpy(),
                                 prob[i, j].cpu().numpy(), thr)
        assert np.array_equal(hist[i, j].cpu().numpy(), want)


def test_c1_size_confusion():
    from mlcomp_b200 import scoring
    n, c = 42000, 10
    rng = np.random.default_rng(1)
    y = rng.integers(0, c, n)
    logits = rng.normal(0, 1, (n, c)).astype(np.float32)
    logits[np.arange(n), y] += 3
    sc = scoring.ClassifyScore(c)
    for lo in range(0, n, 8192):                  # chunks accumulate
        sc.add(torch.from_numpy(logits[lo:lo + 8192]).to(DEV),
               torch.from_numpy(y[lo:lo + 8192]).to(DEV))
    r = sc.result()
    assert np.array_equal(r['confusion'], oracle.confusion(y, logits))
    assert r['accuracy'] == float(oracle.classify_score_final(
        oracle.classify_scores(logits, y)))
    assert r['confusion'].sum() == n and r['dropped'] == 0


def test_device_sigmoid_is_monotone_over_all_fp32():
    """The bounded RLE path for fp32 sources turns act(x) >= thr into
    x >= x'_c; that needs the device sigmoid (the arithmetic every kernel
    shares, here through mlc_seg_prob) to be non-decreasing over ALL 2^32
    fp32 inputs.  Exhaustive sweep, NaNs excluded."""
    from mlcomp_b200 import ops
    chunk = 1 << 27
    bad = 0
    for negative in (False, True):
        last = None
        for b0 in range(0, 0x7F800001, chunk):
            n = min(chunk, 0x7F800001 - b0)
            bits = torch.arange(b0, b0 + n, device=DEV, dtype=torch.int64)
            if negative:
                bits = bits - (1 << 31)       # two's complement of 0x8000.. | b
            x = bits.to(torch.int32).view(torch.float32)
            p = ops.seg_prob([x.view(1, 1, 1, n)],
                             activation='sigmoid').view(-1)
            d = p[1:] - p[:-1]
            bad += int(((d > 0) if negative else (d < 0)).sum())
            if last is not None:
                bad += int((p[0] > last) if negative else (p[0] < last))
            last = p[-1].clone()
            del bits, x, p, d
    assert bad == 0
