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


@pytest.mark.parametrize('c', [2, 3, 5, 6, 7, 8])
@pytest.mark.parametrize('n_src', [1, 2, 3])
def test_argmax_item_kernel_ties_and_near_ties(ops, c, n_src):
    """seg_argmax_items (fp16, rows of whole items, 2..8 classes): logits on a
    coarse grid make exact ties and near-ties frequent, so both the decided
    and the exactly evaluated pixels are exercised; labels must equal the
    arg-max of the device's exact scores, counts and confusion the oracle's."""
    rng = np.random.default_rng(100 * c + n_src)
    n, h, w = 2, 12, 64
    tags = [[], ['hflip'], ['hflip', 'vflip']][:n_src]
    base = rng.integers(-3, 4, (n, c, h, w)).astype(np.float32) * 0.5
    base[0, :, 0, :8] = 0.0                      # all classes equal
    base[0, :, 1, :8] = 11.0                     # ... and saturated
    base[1, 0, 2, 3] = np.float32('nan')
    base[1, c - 1, 2, 4] = np.float32('inf')
    base[1, :, 2, 5] = np.float32('-inf')
    base[1, 1, 3, :] += 2.0 ** -10               # one fp16 ulp at 1.0 or less
    srcs = [torch.from_numpy(base + (rng.integers(0, 2, base.shape) * 2.0 ** -9
                                     if k else 0)).half()
            for k in range(n_src)]
    seen = [forward_tta(s_, t) for s_, t in zip(srcs, tags)]
    dev = [s_.to(DEV) for s_ in seen]
    gt_label = torch.from_numpy(rng.integers(0, c + 1, (n, h, w)).astype(np.uint8))
    for act in ('softmax', None):
        own = ops.seg_prob(dev, xforms=tags, activation=act).cpu().numpy()
        want = np.zeros((n, h, w), np.int64)
        # numpy's order on the device's exact scores: first NaN, else first max
        want = np.argmax(own, axis=1)
        out = ops.seg_argmax(dev, tags, act, gt_label=gt_label.to(DEV),
                             want_onehot=True, want_counts=True,
                             want_confusion=True)
        label = out['label'].cpu().numpy()
        assert np.array_equal(label, want), int((label != want).sum())
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


@pytest.mark.parametrize('c', [2, 5, 8])
def test_argmax_item_kernel_extreme_logits(ops, c):
    """The estimate takes no maximum off the logits: values whose exponentials
    overflow, flush to zero or leave the sum outside [2^-100, 2^128) must reach
    the exact path and give the labels of the exact scores."""
    rng = np.random.default_rng(31 + c)
    n, h, w = 2, 8, 64
    levels = np.array([-120., -100., -88., -87.5, -87.25, -87., -86.5, -70.,
                       -69., -68.5, 0., 60., 86., 87., 88., 88.5, 89., 100.,
                       2000.], np.float32)
    base = levels[rng.integers(0, len(levels), (n, c, h, w))]
    # the same regime for all classes of a pixel in half of the rows, so that
    # near-ties inside one regime occur too
    same = levels[rng.integers(0, len(levels), (n, 1, h // 2, w))]
    base[:, :, :h // 2] = same + rng.integers(-2, 3, (n, c, h // 2, w)) * 0.25
    tags = [[], ['hflip']]
    srcs = [torch.from_numpy(base).half(),
            torch.from_numpy(base + rng.integers(-1, 2, base.shape) * 0.125)
            .half()]
    seen = [forward_tta(s_, t) for s_, t in zip(srcs, tags)]
    dev = [s_.to(DEV) for s_ in seen]
    own = ops.seg_prob(dev, xforms=tags, activation='softmax').cpu().numpy()
    want = np.argmax(own, axis=1)
    label = ops.seg_argmax(dev, tags, 'softmax')['label'].cpu().numpy()
    assert np.array_equal(label, want), int((label != want).sum())
    one = ops.seg_argmax(dev[:1], tags[:1], 'softmax')['label'].cpu().numpy()
    own1 = ops.seg_prob(dev[:1], xforms=tags[:1],
                        activation='softmax').cpu().numpy()
    assert np.array_equal(one, np.argmax(own1, axis=1))


@pytest.mark.parametrize('seed', range(12))
def test_argmax_item_kernel_random_geometry(ops, seed):
    """Random class counts, source counts, flips and plane sizes (rows of
    whole items) through seg_argmax_items, both activations, both forms of
    the truth: labels equal the arg-max of the device's exact scores."""
    rng = np.random.default_rng(500 + seed)
    c = int(rng.integers(2, 9))
    n_src = int(rng.integers(1, 4))
    n, h, w = int(rng.integers(1, 4)), int(rng.integers(1, 20)), \
        8 * int(rng.integers(1, 40))
    flips = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]
    tags = [flips[int(k)] for k in rng.integers(0, 4, n_src)]
    scale = float(rng.choice([0.25, 1.0, 4.0]))
    srcs = [torch.from_numpy((rng.normal(0, scale, (n, c, h, w)))
                             .astype(np.float32)).half() for _ in range(n_src)]
    seen = [forward_tta(s_, t) for s_, t in zip(srcs, tags)]
    dev = [s_.to(DEV) for s_ in seen]
    gt_label = torch.from_numpy(rng.integers(0, c + 1, (n, h, w))
                                .astype(np.uint8))
    g1 = np.stack([(gt_label.numpy() == j) for j in range(c)], axis=1)
    for act in ('softmax', None):
        own = ops.seg_prob(dev, xforms=tags, activation=act).cpu().numpy()
        want = np.argmax(own, axis=1)
        out = ops.seg_argmax(dev, tags, act, gt_label=gt_label.to(DEV),
                             want_counts=True, want_confusion=True)
        label = out['label'].cpu().numpy()
        assert np.array_equal(label, want), int((label != want).sum())
        assert np.array_equal(out['counts'].cpu().numpy(),
                              oracle.onehot_counts(label, g1))
        out2 = ops.seg_argmax(dev, tags, act, gt_onehot=torch.from_numpy(
            g1.astype(np.uint8) * int(rng.integers(1, 256))).to(DEV),
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


def test_iou_sums_match_oracle_formula(ops):
    """mlc_dice_iou_search: dice sums unchanged, IoU sums = the oracle's
    (builder-defined, unpinned) formula on the same integer counts."""
    from test_gpu_parity import synth_seg
    n, c, h, w = 7, 3, 48, 64
    srcs, gt = synth_seg(n, c, h, w, seed=12, empty=0.4)
    thr = oracle.threshold_grid(100)
    areas = [0, 10, 200, 5000]
    hist = ops.seg_hist([srcs[0].to(DEV)], gt.to(DEV), thr)
    d0, m0 = ops.dice_search(hist, areas)
    d1, u1, m1 = ops.dice_iou_search(hist, areas)
    assert torch.equal(d0, d1) and torch.equal(m0, m1)
    prob = ops.seg_prob([srcs[0].to(DEV)]).cpu().numpy()
    want = np.zeros((c, len(thr), len(areas)))
    for i in range(n):
        for j in range(c):
            t, p, ii = oracle.plane_counts(gt[i, j].numpy(), prob[i, j], thr)
            for m, a in enumerate(areas):
                want[j, :, m] += oracle.iou_from_counts(t, p, ii, a)
    np.testing.assert_allclose(u1.cpu().numpy(), want, rtol=1e-12, atol=0)


def test_valid_segment_argmax_executor(ops, tmp_path):
    """``type: valid_segment`` with activation 'softmax' and argmax: score =
    mean dice of the one-hot arg-max masks; csv = their run-length strings."""
    from mlcomp_b200.worker import store
    from mlcomp_b200.worker.executors import Executor
    n, c, h, w = 5, 3, 32, 48
    tags = [[], ['hflip']]
    seen, gt_label = make(n, c, h, w, 2, 99, torch.float32, tags)
    g1 = np.stack([(gt_label.numpy() == j) for j in range(c)], axis=1) \
        .astype(np.uint8)
    cfg = {'executors': {'valid': {
        'type': 'valid_segment', 'logits': [s.numpy() for s in seen],
        'tta': tags, 'targets': g1, 'activation': 'softmax', 'argmax': True,
        'part_size': 2, 'prepare_submit': True}}}
    ex = Executor.from_config(executor='valid', config=cfg,
                              additional_info={})
    out = ex(task=store.Task(name='t'), task_provider=None,
             dag=store.Dag(name='d'))
    label = ops.seg_argmax([s.to(DEV) for s in seen], tags, 'softmax')[
        'label'].cpu().numpy()
    dice = np.zeros((n, c))
    for i in range(n):
        for j in range(c):
            m = (label[i] == j).astype(np.float32)
            dice[i, j] = oracle.plane_dice(g1[i, j].astype(np.float32), m,
                                           threshold=0.5)
            assert ex.rows[i * c + j]['EncodedPixels'] == \
                oracle.rle_encode(m.astype(np.uint8))
    assert abs(out['res'] - dice.mean()) < 1e-12
    np.testing.assert_allclose(out['dice_per_class'], dice.mean(axis=0),
                               rtol=1e-12)
