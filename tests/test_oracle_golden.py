"""Pin oracle/ against outputs of the reference's own functions
(tests/golden/make_golden.py ran them in the build container)."""
import json
import os

import numpy as np
import torch

import oracle


def _planes(golden_dir):
    z = np.load(os.path.join(golden_dir, 'dice_planes.npz'))
    rows = json.load(open(os.path.join(golden_dir, 'dice_rows.json')))
    shapes = {}
    for r in rows:
        shapes[r['case']] = (r['h'], r['w'])
    off = 0
    planes = {}
    for case in sorted(shapes):
        h, w = shapes[case]
        planes[case] = (z['gt'][off:off + h * w].reshape(h, w),
                        z['prob'][off:off + h * w].reshape(h, w))
        off += h * w
    return planes, rows, z['thr']


def test_dice_point_matches_reference(golden_dir):
    planes, rows, thr = _planes(golden_dir)
    assert np.array_equal(thr, oracle.threshold_grid(100))
    for r in rows:
        gt, prob = planes[r['case']]
        if r['k'] < 0:
            got = oracle.plane_dice(gt, (prob >= np.float32(0.5)).astype(np.uint8))
        else:
            got = oracle.plane_dice(gt, prob, threshold=np.float32(thr[r['k']]),
                                    min_area=r['min_area'],
                                    empty_one=r['empty_one'])
        assert isinstance(got, int) == r['is_int'], r
        assert float(got) == r['value'], r          # bit-equal fp64


def test_dice_from_integer_counts_is_bit_equal(golden_dir):
    planes, rows, thr = _planes(golden_dir)
    cache = {}
    for r in rows:
        if r['k'] < 0:
            continue
        c = r['case']
        if c not in cache:
            cache[c] = oracle.plane_counts(*planes[c], thr)
        t, p, i = cache[c]
        got = oracle.dice_from_counts(t, p[r['k']], i[r['k']], r['min_area'],
                                      r['empty_one'])
        assert float(got) == r['value'], r


def test_rle_matches_reference(golden_dir):
    z = np.load(os.path.join(golden_dir, 'rle_masks.npz'))
    rows = json.load(open(os.path.join(golden_dir, 'rle_rows.json')))
    off = 0
    for r in rows:
        h, w = r['h'], r['w']
        m = z['mask'][off:off + h * w].reshape(h, w)
        off += h * w
        assert oracle.rle_encode(m) == r['rle']
        assert np.array_equal(oracle.rle_decode(r['rle'], (w, h)), m)
    assert oracle.rle_encode(np.zeros((3, 4), np.uint8)) == ''
    assert oracle.rle_encode(np.ones((3, 4), np.uint8)) == '1 12'


def test_flip_matches_reference(golden_dir):
    z = np.load(os.path.join(golden_dir, 'flip.npz'))
    x = torch.from_numpy(z['x'])
    assert np.array_equal(oracle.flip_axis(x, 3).numpy(), z['last'])
    assert np.array_equal(oracle.flip_axis(x, 2).numpy(), z['prev'])
    assert np.array_equal(oracle.tta_inverse(x, ['hflip', 'vflip']).numpy(),
                          z['both'])
    assert np.array_equal(oracle.tta_inverse(x, ['transpose']).numpy(),
                          z['x'].transpose(0, 1, 3, 2))


def test_confusion_matches_reference(golden_dir):
    z = np.load(os.path.join(golden_dir, 'confusion.npz'))
    assert np.array_equal(oracle.confusion(z['y'], z['preds']), z['cm'])
    assert np.array_equal(oracle.confusion_numpy(z['y'], z['preds']), z['cm'])
    assert oracle.confusion(z['y'], z['preds']).dtype == np.int64


def test_grid_slow_equals_fast():
    rng = np.random.default_rng(3)
    gt = (rng.random((3, 2, 9, 11)) < 0.3).astype(np.uint8)
    gt[0, 0] = 0
    prob = rng.random((3, 2, 9, 11)).astype(np.float32)
    thr = oracle.threshold_grid(20)
    areas = [0, 3, 30, 90]
    slow = oracle.dice_grid_slow(gt, prob, thr, areas)
    fast = oracle.dice_grid_fast(gt, prob, thr, areas)
    assert np.array_equal(slow, fast)


def test_activation_errors_like_reference():
    import pytest
    with pytest.raises(Exception, match='unknown activation'):
        oracle.activate(torch.zeros(2, 2), 'relu')
    with pytest.raises(Exception, match='Unknwon augmentation'):
        oracle.tta_inverse(torch.zeros(1, 1, 2, 2), ['rot90'])


def test_composites_match_reference(golden_dir):
    """encode_pred / plot_mask (segmenation.py:71-96) as run by
    make_golden_composite.py with OpenCV 4.13."""
    z = np.load(os.path.join(golden_dir, 'composite.npz'))
    for k in range(int(z['n_cases'])):
        colors = [tuple(int(v) for v in c) for c in z[f'colors_{k}']] or None
        mask, img = z[f'mask_{k}'], z[f'img_{k}']
        assert np.array_equal(oracle.encode_pred(mask, colors),
                              z[f'encode_{k}']), k
        assert np.array_equal(oracle.plot_mask(img, mask, colors),
                              z[f'plot_{k}']), k
        assert np.array_equal(oracle.plot_mask(img[:, :, 0], mask, colors),
                              z[f'plot_gray_{k}']), k
