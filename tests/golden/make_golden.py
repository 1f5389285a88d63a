"""Generate tests/golden/*.npz|json by running the REFERENCE's own functions.

Run in the build container only (``/root/reference`` does not exist on the
GPU box):  ``python tests/golden/make_golden.py``

The reference leaves are loaded BY FILE PATH - ``import mlcomp`` has side
effects (creates ~/mlcomp, copies .env; mlcomp/__init__.py:7-51) and most of
the package needs dependencies that are not installed (albumentations,
sqlalchemy).  Loaded files:
    mlcomp/contrib/metrics/dice.py     -> dice_numpy
    mlcomp/contrib/transform/rle.py    -> mask2rle, rle2mask
    mlcomp/contrib/torch/tensors.py    -> flip
and the third-party call of mlcomp/worker/reports/classification.py:133-138
(sklearn.metrics.confusion_matrix) is issued verbatim.
The fixtures pin ``oracle/`` (tests/test_oracle_golden.py); the oracle then
checks the CUDA path.
"""
import importlib.util
import json
import os

import numpy as np
import torch

REF = '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))


def load(rel, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def blobs(rng, h, w, n):
    m = np.zeros((h, w), np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n):
        cy, cx = rng.integers(0, h), rng.integers(0, w)
        ry, rx = rng.integers(1, max(2, h // 3)), rng.integers(1, max(2, w // 3))
        m |= (((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 <= 1).astype(np.uint8)
    return m


def main():
    dice = load('mlcomp/contrib/metrics/dice.py', 'ref_dice').dice_numpy
    rle = load('mlcomp/contrib/transform/rle.py', 'ref_rle')
    flip = load('mlcomp/contrib/torch/tensors.py', 'ref_tensors').flip
    from sklearn.metrics import confusion_matrix

    rng = np.random.default_rng(20260311)
    thr = (np.arange(100) / 100).astype(np.float32)

    # ---------------- dice_numpy ------------------------------------------
    gts, probs, rows = [], [], []
    shapes = [(16, 24), (7, 13), (32, 32), (1, 40), (40, 1)]
    for case in range(40):
        h, w = shapes[case % len(shapes)]
        kind = case % 8
        if kind == 0:
            gt = np.zeros((h, w), np.uint8)
        elif kind == 1:
            gt = np.ones((h, w), np.uint8)
        else:
            gt = blobs(rng, h, w, rng.integers(1, 4))
        logit = 6 * (gt.astype(np.float32) - 0.5) + rng.normal(0, 2.5, (h, w))
        prob = torch.sigmoid(torch.from_numpy(logit.astype(np.float32))).numpy()
        if kind == 2:                    # probabilities exactly on the grid
            prob = rng.choice(thr, size=(h, w)).astype(np.float32)
        if kind == 3:                    # +-1 ulp neighbours of grid points
            base = rng.choice(thr[1:], size=(h, w)).astype(np.float32)
            step = rng.integers(-1, 2, size=(h, w))
            prob = np.where(step > 0, np.nextafter(base, np.float32(2)),
                            np.where(step < 0,
                                     np.nextafter(base, np.float32(-1)), base)
                            ).astype(np.float32)
        if kind == 4:
            prob = np.zeros((h, w), np.float32)      # nothing predicted > 0
        gts.append(gt.astype(np.float32).ravel())
        probs.append(prob.ravel())
        gt_f = gt.astype(np.float32)     # contrib/dataset/segment.py:36-56
        for k in (0, 1, 17, 50, 73, 99):
            for area in (None, 0, 1, 5, 40, 10 ** 6):
                for empty_one in (True, False):
                    r = dice(gt_f, prob, threshold=np.float32(thr[k]),
                             min_area=area, empty_one=empty_one)
                    rows.append(dict(case=case, h=h, w=w, k=k,
                                     min_area=area, empty_one=empty_one,
                                     is_int=isinstance(r, int),
                                     value=float(r)))
        # threshold=None: outputs already binary
        pred = (prob >= np.float32(0.5)).astype(np.uint8)
        r = dice(gt_f, pred)
        rows.append(dict(case=case, h=h, w=w, k=-1, min_area=None,
                         empty_one=True, is_int=isinstance(r, int),
                         value=float(r)))
    np.savez_compressed(os.path.join(HERE, 'dice_planes.npz'),
                        gt=np.concatenate(gts), prob=np.concatenate(probs),
                        thr=thr)
    json.dump(rows, open(os.path.join(HERE, 'dice_rows.json'), 'w'))

    # ---------------- mask2rle / rle2mask ---------------------------------
    rle_rows = []
    masks = []
    for case in range(24):
        h, w = [(8, 8), (5, 9), (32, 7), (33, 65), (1, 17), (17, 1),
                (64, 64), (2, 2)][case % 8]
        kind = case % 6
        if kind == 0:
            m = np.zeros((h, w), np.uint8)
        elif kind == 1:
            m = np.ones((h, w), np.uint8)
        elif kind == 2:
            m = (rng.random((h, w)) < 0.5).astype(np.uint8)   # 1-px runs
        elif kind == 3:                 # runs spanning column ends
            m = np.zeros((h, w), np.uint8)
            m[h // 2:, :] = 1
            m[:1, 1:] = 1
        else:
            m = blobs(rng, h, w, 2)
        s = rle.mask2rle(m)
        back = rle.rle2mask(s, (w, h))
        assert back.shape == (h, w) and (back == m).all()
        masks.append(m.ravel())
        rle_rows.append(dict(h=h, w=w, rle=s))
    np.savez_compressed(os.path.join(HERE, 'rle_masks.npz'),
                        mask=np.concatenate(masks))
    json.dump(rle_rows, open(os.path.join(HERE, 'rle_rows.json'), 'w'))

    # ---------------- flip -------------------------------------------------
    x = torch.arange(2 * 3 * 4 * 5, dtype=torch.float32).reshape(2, 3, 4, 5)
    np.savez_compressed(os.path.join(HERE, 'flip.npz'), x=x.numpy(),
                        last=flip(x, 3).numpy(), prev=flip(x, 2).numpy(),
                        both=flip(flip(x, 3), 2).numpy())

    # ---------------- confusion matrix ------------------------------------
    n, c = 500, 10
    y = rng.integers(0, c, n)
    preds = rng.normal(0, 1, (n, c)).astype(np.float32)
    preds[np.arange(n), y] += 1.5
    preds[:20, 3] = preds[:20, 7] = 9.0            # arg-max ties -> lowest
    cm = np.array(confusion_matrix(y, preds.argmax(axis=1),
                                   labels=np.arange(preds.shape[1])))
    np.savez_compressed(os.path.join(HERE, 'confusion.npz'), y=y, preds=preds,
                        cm=cm)
    print('golden fixtures written to', HERE)


if __name__ == '__main__':
    main()
