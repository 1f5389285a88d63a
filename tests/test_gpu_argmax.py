"""GPU parity of the segmentation arg-max path (mlc_seg_argmax): label maps,
one-hot masks, their {T, P, I} counts, the pixel confusion matrix and the RLE
of the one-hot masks, bit-exact against the oracle
(np.argmax(oracle.fuse_probs(..., 'softmax'), axis=1); reference lines
utils/torch.py:16-17, worker/reports/classification.py:111,136,
contrib/metrics/dice.py:10-11,20, contrib/transform/rle.py:4-13)."""
import numpy as np
import pytest
import torch

import oracle
from test_gpu_parity import DEV, forward_tta

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def ops():
    from mlcomp_b200 import ops as o
    return o


def make(n, c, h, w, n_src, seed, dtype, tags):
    rng = np.random.default_rng(seed)
    gt_label = rng.integers(0, c + 1, (n, h, w)).astype(np.uint8)  # c = none
    gt_label[gt_label == c] = 255
    base = rng.normal(0, 1, (n, c, h, w)).astype(np.float32)
    for j in range(c):
        base[:, j][gt_label == j] += 2.0
    srcs = [torch.from_numpy(base + rng.normal(0, .3, base.shape)
                             .astype(np.float32)).to(dtype)
            for _ in range(n_src)]
    seen = [forward_tta(s, t) for s, t in zip(srcs, tags)]
    return seen, torch.from_numpy(gt_label)


@pytest.mark.parametrize('dtype', [torch.float16, torch.float32])
@pytest.mark.parametrize('c', [2, 4, 5, 21])
@pytest.mark.parametrize('shape,tags', [
    ((2, 35, 53), [[]]),
    ((2, 32, 40), [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]),
    ((1, 24, 24), [[], ['transpose'], ['hflip', 'transpose']]),
])
def test_argmax_matches_oracle(ops, dtype, c, shape, tags):
    n, h, w = shape
    seen, gt_label = make(n, c, h, w, len(tags), 7 * c + h, dtype, tags)
    dev = [s.to(DEV) for s in seen]
    for act in ('softmax', None):
        want = oracle.seg_argmax(seen, tags, act)
        # the device's own scores give the same labels (exact arithmetic match)
        own = ops.seg_prob(dev, xforms=tags, activation=act).cpu().numpy()
        out = ops.seg_argmax(dev, tags, act, gt_label=gt_label.to(DEV),
                             want_onehot=True, want_counts=True,
                             want_confusion=True)
        label = out['label'].cpu().numpy()
        assert np.array_equal(label, np.argmax(own, axis=1))
        # against the CPU oracle: only pixels whose two best scores are within
        # a few ulp (CPU vs GPU exp) may differ
        diff = label != want
        if diff.any():
            cpu = oracle.fuse_probs(seen, tags, act)
            top2 = np.sort(cpu, axis=1)[:, -2:]
            gap = (top2[:, 1] - top2[:, 0])[diff]
            assert (gap <= 1e-6).all()
            assert diff.sum() <= 2 + label.size // 20000
        onehot = out['onehot'].cpu().numpy()
        for j in range(c):
            assert np.array_equal(onehot[:, j], (label == j).astype(np.uint8))
        g1 = np.stack([(gt_label.numpy() == j) for j in range(c)], axis=1)
        assert np.array_equal(out['counts'].cpu().numpy(),
                              oracle.onehot_counts(label, g1))
        keep = gt_label.numpy() < c
        cm = np.bincount(gt_label.numpy()[keep].astype(np.int64) * c +
                         label[keep], minlength=c * c).reshape(c, c)
        assert np.array_equal(out['confusion'].cpu().numpy(), cm)
        # one-hot truth gives the same counts
        out2 = ops.seg_argmax(dev, tags, act,
                              gt_onehot=torch.from_numpy(
                                  g1.astype(np.uint8)).to(DEV),
                              want_label=False, want_counts=True)
        assert torch.equal(out2['counts'], out['counts'])


def test_argmax_nan_and_ties(ops):
    """numpy order: first NaN wins, else first maximum."""
    x = torch.zeros((1, 4, 1, 8), dtype=torch.float32)
    x[0, :, 0, 0] = torch.tensor([1., 3., 3., 2.])          # tie -> 1
    x[0, :, 0, 1] = torch.tensor([1., float('nan'), 5., float('nan')])  # 1
    x[0, :, 0, 2] = torch.tensor([float('nan'), 0., 0., 0.])  # 0
    x[0, :, 0, 3] = torch.tensor([0., 0., 0., 0.])          # all equal -> 0
    x[0, :, 0, 4] = torch.tensor([-1., -2., float('inf'), float('inf')])  # 2
    x[0, :, 0, 5] = torch.tensor([float('-inf')] * 4)       # 0
    x[0, :, 0, 6] = torch.tensor([0., 1e-30, 0., 0.])       # 1
    x[0, :, 0, 7] = torch.tensor([5., 4., 3., float('nan')])  # 3
    out = ops.seg_argmax([x.to(DEV)], None, None)
    want = np.argmax(x.numpy(), axis=1)
    assert np.array_equal(out['label'].cpu().numpy(), want)
    assert want[0, 0].tolist() == [1, 1, 0, 0, 2, 0, 1, 3]
    for dtype in (torch.float16, torch.float32):
        o2 = ops.seg_argmax([x.to(dtype).to(DEV)], None, 'softmax')
        p = ops.seg_prob([x.to(dtype).to(DEV)], None, 'softmax').cpu().numpy()
        assert np.array_equal(o2['label'].cpu().numpy(), np.argmax(p, axis=1))


def test_argmax_severstal_shape_rle_and_dice(ops):
    """256x1600, 4 classes + background, 2 flips: RLE of the one-hot masks
    equals mask2rle of (label == c); dice of the one-hot masks from the
    counts equals dice_numpy."""
    n, c, h, w = 2, 5, 256, 1600
    tags = [[], ['hflip']]
    seen, gt_label = make(n, c, h, w, 2, 3, torch.float16, tags)
    dev = [s.to(DEV) for s in seen]
    out = ops.seg_argmax(dev, tags, 'softmax', gt_label=gt_label.to(DEV),
                         want_onehot=True, want_counts=True)
    label = out['label'].cpu().numpy()
    runs, off, cnt, area = ops.rle_encode([out['onehot']], 0.5, 0,
                                          activation=None)
    strings = ops.rle_strings(runs, off, cnt)
    counts = out['counts'].cpu().numpy()
    for i in range(n):
        for j in range(c):
            m = (label[i] == j).astype(np.uint8)
            assert strings[i * c + j] == oracle.rle_encode(m)
            g = (gt_label[i].numpy() == j).astype(np.float32)
            t, p, ii = counts[i, j]
            assert oracle.dice_from_counts(t, p, ii) == \
                oracle.plane_dice(g, m.astype(np.float32), threshold=0.5)
