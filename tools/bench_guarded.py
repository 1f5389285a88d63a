#!/usr/bin/env python
"""Kernel timings of the multi-source fp16 histogram pass (seg_hist_guarded
vs the table kernel it replaces), for A/B runs over library variants:

    MLC_LIB=mlcomp_b200/lib/libmlcomp_b200_dev.so MLC_HIST_VARIANT=6 \
        python tools/bench_guarded.py tta c5      # the exact table kernel
    python tools/bench_guarded.py tta c5 rle     # default library

One JSON object per line; `tag` names the library / variant.
"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import CFG, TAGS, synth_tta  # noqa: E402
from mlcomp_b200 import _lib, ops  # noqa: E402

DEV = torch.device('cuda', 0)
PEAK = 6553.3
try:
    PEAK = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    pass
TAG = os.path.basename(os.environ.get('MLC_LIB', 'default')) + \
    (':V' + os.environ['MLC_HIST_VARIANT']
     if 'MLC_HIST_VARIANT' in os.environ else '')


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(iters)]
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


def slow_px():
    lib = _lib.load()
    if not hasattr(lib, 'mlc_dev_guard_slow_pixels'):
        return None
    fn = lib.mlc_dev_guard_slow_pixels
    fn.restype = ctypes.c_longlong
    return int(fn())


def line(name, ms, nbytes, planes, **extra):
    gbs = nbytes / (ms * 1e-3) / 1e9
    out = {'tag': TAG, 'case': name, 'ms': round(ms, 4), 'GB/s': round(gbs, 1),
           'frac_of_measured_peak': round(gbs / PEAK, 3),
           'masks_per_s': round(planes / (ms * 1e-3))}
    out.update(extra)
    print(json.dumps(out), flush=True)


def tta(n=512, sharp=False):
    cfg = dict(CFG)
    srcs, gt = synth_tta(cfg, n, 7, DEV)
    if sharp:   # confident logits: N(+-6, 1)
        srcs = [(s.float() * 0.66 + (s.float() > 0) * 4 - 2).half()
                for s in srcs]
    thr = ops.threshold_grid(100)
    px = n * 4 * 256 * 1600
    name = 'tta4_%s_%dimg' % ('sharp' if sharp else 'synth', n)
    slow_px()
    ops.seg_hist(srcs, gt, thr, xforms=TAGS)
    sp = slow_px()
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=TAGS))
    line(name + '_hist', ms, px * 9 + n * 4 * 808, n * 4,
         exact_path_pixel_fraction=None if sp is None else sp / px)
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=TAGS, bins=True))
    line(name + '_hist+bins', ms, px * 9 + n * 4 * 808, n * 4)
    return srcs, gt, thr


def c5(n=512):
    cfg = dict(CFG)
    srcs, gt = synth_tta(cfg, n, 11, DEV, n_var=5)
    # ensemble members are not flipped: undo the storage flips of synth_tta
    from bench import FLIP_DIMS
    srcs = [s.flip(*FLIP_DIMS[k % 4]).contiguous() if FLIP_DIMS[k % 4] else s
            for k, s in enumerate(srcs)]
    thr = ops.threshold_grid(100)
    px = n * 4 * 256 * 1600
    slow_px()
    ops.seg_hist(srcs, gt, thr)
    sp = slow_px()
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr))
    line('c5_5models_%dimg_hist' % n, ms, px * 11 + n * 4 * 808, n * 4,
         exact_path_pixel_fraction=None if sp is None else sp / px)


def rle(n=512):
    """Pipeline tail: RLE from the bin planes at the point a search picks."""
    from mlcomp_b200 import scoring
    cfg = dict(CFG)
    srcs, gt = synth_tta(cfg, n, 7, DEV)
    thr = ops.threshold_grid(100)
    s = scoring.SegSearch(4, thr, cfg['min_areas'], device=DEV)
    s.reset()
    hist, bins = s.add(srcs, gt, xforms=TAGS, keep_bins=True)
    best = s.result()['best']
    k_c = [b['threshold_index'] for b in best]
    a_c = [b['min_area'] for b in best]
    t_c = [b['threshold'] for b in best]
    px = n * 4 * 256 * 1600
    cap = px // 32
    r = ops.rle_from_bins(bins, k_c, a_c, capacity=cap, hist=hist)
    nonempty = int((r[2] > 0).sum())
    runs = int(r[0].numel() // 2)
    ms = timeit(lambda: ops.rle_from_bins(bins, k_c, a_c, capacity=cap,
                                          hist=hist))
    line('rle_bins_known_%dimg' % n, ms, px + 8 * runs, n * 4,
         nonempty_masks=nonempty, runs=runs, best=[k_c, a_c])
    ms = timeit(lambda: ops.rle_from_bins(bins, k_c, a_c, capacity=cap))
    line('rle_bins_all_%dimg' % n, ms, px + 8 * runs, n * 4)
    area = ops.plane_area(hist, k_c)
    ms = timeit(lambda: ops.rle_encode(srcs, t_c, a_c, xforms=TAGS,
                                       capacity=cap, known_area=area))
    line('rle_sources_known_%dimg' % n, ms, px * 8 + 8 * runs, n * 4)
    ms = timeit(lambda: ops.rle_encode(srcs, t_c, a_c, xforms=TAGS,
                                       capacity=cap))
    line('rle_sources_all_%dimg' % n, ms, px * 8 + 8 * runs, n * 4)


if __name__ == '__main__':
    what = sys.argv[1:] or ['tta', 'c5']
    for wname in what:
        if wname == 'tta':
            tta()
        elif wname == 'sharp':
            tta(sharp=True)
        elif wname == 'c5':
            c5()
        elif wname == 'rle':
            rle()
