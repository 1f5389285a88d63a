#!/usr/bin/env python
"""Randomised parity sweep on a GPU box (not part of the pytest suite):
random shapes, dtypes, source counts, transforms, thresholds and tensor
offsets for seg_prob / seg_hist / rle_encode / rle_decode against the oracle.

    python tools/fuzz_gpu.py [n_cases] [seed]
"""
import itertools
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from mlcomp_b200 import ops  # noqa: E402

DEV = 'cuda:0'


def main():
    n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    tag_sets = [list(t) for k in range(3)
                for t in itertools.product(['hflip', 'vflip'], repeat=k)]
    bad = 0
    for case in range(n_cases):
        n, c = int(rng.integers(1, 4)), int(rng.integers(1, 5))
        big = os.environ.get('FUZZ_BIG') == '1'   # several bands / x-chunks
        h = int(rng.integers(1, 300 if big else 70))
        w = int(rng.integers(1, 700 if big else 130))
        if rng.random() < 0.3:
            w = int(rng.integers(1, 80 if big else 9)) * 8
        n_src = int(rng.integers(1, 6))
        dtype = torch.float16 if rng.random() < 0.7 else torch.float32
        act = ['sigmoid', None, 'softmax'][int(rng.integers(0, 3))]
        tags = [tag_sets[int(rng.integers(0, len(tag_sets)))]
                for _ in range(n_src)]
        if rng.random() < 0.15 and h != w:
            tags[0] = tags[0] + ['transpose']
        gt = (rng.random((n, c, h, w)) < rng.random()).astype(np.uint8)
        base = 6 * (gt.astype(np.float32) - 0.5) + \
            rng.normal(0, 2.5, gt.shape).astype(np.float32)
        srcs = []
        for t in tags:
            x = torch.from_numpy(base + rng.normal(0, 1, gt.shape)
                                 .astype(np.float32)).to(dtype)
            if act is None:
                x = torch.sigmoid(x.float()).to(dtype)
            srcs.append(oracle.tta_inverse(x, list(reversed(t))).contiguous())
        # put every source at an odd offset inside a larger buffer half of
        # the time: exercises the unaligned window loads
        dev_src = []
        for s in srcs:
            if rng.random() < 0.5:
                pad = int(rng.integers(1, 9))
                buf = torch.zeros(s.numel() + pad, dtype=dtype, device=DEV)
                view = buf[pad:].view(s.shape)
                view.copy_(s)
                dev_src.append(view)
            else:
                dev_src.append(s.to(DEV))
        b = int(rng.integers(1, 120))
        thr = oracle.threshold_grid(b) if rng.random() < 0.7 else \
            np.sort(rng.random(b)).astype(np.float32)
        try:
            prob = ops.seg_prob(dev_src, xforms=tags, activation=act)
            want = oracle.fuse_probs(srcs, tags, act)
            p = prob.cpu().numpy()
            ulps = np.abs(p.view(np.int32).astype(np.int64) -
                          want.view(np.int32).astype(np.int64))
            lim = 0 if act is None else 16 * n_src
            assert ulps.max() <= lim, ('prob', ulps.max())
            lut = ops.seg_lut(thr, act) if act != 'softmax' else None
            hist = ops.seg_hist(dev_src, torch.from_numpy(gt).to(DEV), thr,
                                xforms=tags, activation=act, lut=lut)
            hh = hist.cpu().numpy()
            for i in range(n):
                for j in range(c):
                    assert np.array_equal(
                        hh[i, j], oracle.bin_counts(gt[i, j], p[i, j], thr)), \
                        ('hist', i, j)
            if b <= 255:
                h2, bins = ops.seg_hist(dev_src, torch.from_numpy(gt).to(DEV),
                                        thr, xforms=tags, activation=act,
                                        bins=True)
                assert torch.equal(h2, hist), ('hist with bins',)
                with np.errstate(invalid='ignore'):
                    wb = np.where(np.isnan(p), 0, np.searchsorted(
                        thr, np.nan_to_num(p), side='right'))
                assert np.array_equal(bins.cpu().numpy(), wb), ('bins',)
                kk = rng.integers(0, b, c)
                ra = ops.rle_from_bins(bins, kk, 3)
                rb = ops.rle_encode(dev_src, thr[kk], 3, xforms=tags,
                                    activation=act)
                assert ops.rle_strings(*ra[:3]) == ops.rle_strings(*rb[:3]), \
                    ('rle from bins',)
            t_c = rng.random(c).astype(np.float32)
            a_c = rng.integers(0, 40, c).astype(np.int32)
            runs, off, cnt, area = ops.rle_encode(dev_src, t_c, a_c,
                                                  xforms=tags, activation=act)
            strs = ops.rle_strings(runs, off, cnt)
            k = 0
            for i in range(n):
                for j in range(c):
                    assert strs[k] == oracle.rle_submission(
                        p[i, j], t_c[j], int(a_c[j])), ('rle', i, j)
                    k += 1
            back = ops.rle_decode(runs, off.reshape(-1), cnt.reshape(-1), h, w)
            k = 0
            for i in range(n):
                for j in range(c):
                    m = (p[i, j] >= t_c[j])
                    if a_c[j] and m.sum() < a_c[j]:
                        m[:] = False
                    assert np.array_equal(back[k].cpu().numpy(),
                                          m.astype(np.uint8)), ('dec', i, j)
                    k += 1
        except AssertionError as e:
            bad += 1
            print('FAIL case', case, dict(n=n, c=c, h=h, w=w, n_src=n_src,
                                          dtype=str(dtype), act=act, tags=tags,
                                          b=b), e.args, flush=True)
    print(f'{n_cases - bad} of {n_cases} random cases passed')
    sys.exit(1 if bad else 0)


if __name__ == '__main__':
    main()
