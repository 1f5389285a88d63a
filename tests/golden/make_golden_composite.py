"""Generate tests/golden/composite.npz by running the REFERENCE's own
``SegmentationReportBuilder.encode_pred`` and ``.plot_mask``
(mlcomp/worker/reports/segmenation.py:71-96).

Run in the build container only:  ``python tests/golden/make_golden_composite.py``

The module cannot be imported (its top level needs sqlalchemy), so the two
methods are cut out of the reference file with ``ast`` AT GENERATION TIME and
executed unchanged, with numpy and this container's OpenCV (4.13) in scope,
on a stand-in object that only carries ``colors``.  Nothing of the reference
is stored in the repo; the fixture holds inputs and outputs only.
"""
import ast
import os

import cv2
import numpy as np

REF = '/root/reference/mlcomp/worker/reports/segmenation.py'
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_methods():
    tree = ast.parse(open(REF).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef)
               and n.name == 'SegmentationReportBuilder')
    keep = [n for n in cls.body if isinstance(n, ast.FunctionDef)
            and n.name in ('encode_pred', 'plot_mask')]
    assert len(keep) == 2
    mod = ast.Module(body=[ast.ClassDef(
        name='Cut', bases=[], keywords=[], body=keep, decorator_list=[],
        type_params=[])], type_ignores=[])
    ast.fix_missing_locations(mod)
    ns = {'np': np, 'cv2': cv2}
    exec(compile(mod, REF, 'exec'), ns)
    return ns['Cut']


def blobs(rng, h, w, n):
    m = np.zeros((h, w), np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n):
        cy, cx = rng.integers(0, h), rng.integers(0, w)
        ry, rx = rng.integers(1, max(2, h // 3)), rng.integers(1, max(2, w // 3))
        m |= (((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 <= 1).astype(np.uint8)
    return m


def main():
    Cut = reference_methods()
    rng = np.random.default_rng(20260312)
    out = {}
    cases = [
        # (C, H, W, kind, colours)
        (1, 12, 17, 'blobs', None),
        (3, 24, 40, 'blobs', [(255, 0, 0), (0, 255, 0), (10, 20, 250)]),
        (4, 33, 29, 'noise', [(200, 100, 90), (100, 200, 90), (90, 90, 200),
                              (255, 255, 255)]),
        (2, 8, 8, 'full', None),
        (2, 9, 11, 'empty', [(1, 2, 3), (4, 5, 6)]),
        (3, 40, 64, 'holes', [(255, 255, 0), (0, 255, 255), (255, 0, 255)]),
        (2, 1, 30, 'noise', None),
        (2, 30, 1, 'noise', None),
        (5, 16, 16, 'noise', [(250, 250, 250)] * 5),     # uint8 sums wrap
    ]
    for k, (c, h, w, kind, colors) in enumerate(cases):
        if kind == 'blobs':
            mask = np.stack([blobs(rng, h, w, 2) for _ in range(c)])
        elif kind == 'noise':
            mask = (rng.random((c, h, w)) < 0.4).astype(np.uint8)
        elif kind == 'full':
            mask = np.ones((c, h, w), np.uint8)
        elif kind == 'empty':
            mask = np.zeros((c, h, w), np.uint8)
        else:   # blobs with holes punched in
            mask = np.stack([blobs(rng, h, w, 3) for _ in range(c)])
            mask &= (rng.random((c, h, w)) < 0.93).astype(np.uint8)
        img = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
        obj = Cut()
        obj.colors = colors
        enc = obj.encode_pred(mask)
        plot = obj.plot_mask(img.copy(), mask)
        gray = obj.plot_mask(img[:, :, 0].copy(), mask)   # 2-D image branch
        out[f'mask_{k}'] = mask
        out[f'img_{k}'] = img
        out[f'colors_{k}'] = np.asarray(colors if colors is not None else [],
                                        np.int64).reshape(-1, 3)
        out[f'encode_{k}'] = enc
        out[f'plot_{k}'] = plot
        out[f'plot_gray_{k}'] = gray
    out['n_cases'] = np.int64(len(cases))
    np.savez_compressed(os.path.join(HERE, 'composite.npz'), **out)
    print('composite.npz:', len(cases), 'cases, OpenCV', cv2.__version__)


if __name__ == '__main__':
    main()
