#!/usr/bin/env python
"""Kernel-level timings for the BASELINE.json configs other than the headline
one (those are parity-test cases for bench.py; this script gives the numbers
DESIGN.md quotes for them).  FULL image counts by default (C3 5,546 images,
C4 10,000, C5 3,750 = one GPU's share of 30,000 over 8), "name:count" for a
reduced one; inputs are far larger than L2.  One JSON object per line.

    python tools/bench_configs.py [c1 c3 c4 c5 c2g c3:1024 ...]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlcomp_b200 import ops  # noqa: E402

DEV = torch.device('cuda', 0)
PEAK = 6550.4
try:
    PEAK = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    pass


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


def synth(n, c, h, w, n_src, seed, empty):
    g = torch.Generator(device=DEV).manual_seed(seed)
    gt = torch.zeros((n, c, h, w), dtype=torch.uint8, device=DEV)
    yy = torch.arange(h, device=DEV, dtype=torch.float32)[None, :, None]
    xx = torch.arange(w, device=DEV, dtype=torch.float32)[None, None, :]
    p = n * c
    cy = torch.rand(p, generator=g, device=DEV) * h
    cx = torch.rand(p, generator=g, device=DEV) * w
    r = (0.05 + 0.2 * torch.rand(p, generator=g, device=DEV)) * min(h, w)
    keep = torch.rand(p, generator=g, device=DEV) >= empty
    for lo in range(0, p, 256):
        hi = min(p, lo + 256)
        m = (((yy - cy[lo:hi, None, None]) / r[lo:hi, None, None]) ** 2 +
             ((xx - cx[lo:hi, None, None]) / (2 * r[lo:hi, None, None])) ** 2
             ) <= 1
        m &= keep[lo:hi, None, None]
        gt.view(p, h, w)[lo:hi] = m.to(torch.uint8)
    srcs = []
    for s in range(n_src):
        x = torch.empty((n, c, h, w), dtype=torch.float16, device=DEV)
        for lo in range(0, n, 64):
            hi = min(n, lo + 64)
            noise = torch.randn((hi - lo, c, h, w), generator=g, device=DEV)
            x[lo:hi] = (gt[lo:hi].float() * 8 - 4 + 1.6 * noise).half()
        srcs.append(x)
    return srcs, gt


RESULTS = []   # every line reported (bench.py collects them as a side key)
QUIET = False


def report(name, ms, nbytes, planes, extra=None):
    gbs = nbytes / (ms * 1e-3) / 1e9
    out = {'config': name, 'ms': round(ms, 4), 'GB/s': round(gbs, 1),
           'frac_of_measured_peak': round(gbs / PEAK, 3),
           'masks_per_s': round(planes / (ms * 1e-3))}
    out.update(extra or {})
    RESULTS.append(out)
    if not QUIET:
        print(json.dumps(out), flush=True)


def c1():
    n, c = 42000, 10
    g = torch.Generator(device=DEV).manual_seed(1)
    y = torch.randint(0, c, (n,), generator=g, device=DEV)
    x = torch.randn((n, c), generator=g, device=DEV)
    x[torch.arange(n), y] += 3
    cm = torch.zeros((c, c), dtype=torch.int64, device=DEV)
    st = torch.zeros(2, dtype=torch.int64, device=DEV)
    ms = timeit(lambda: ops.confusion(x, y, cm=cm, stats=st), iters=50)
    report('C1 confusion 42000x10 fp32 (launch-latency-bound)', ms,
           n * (c * 4 + 8), n, {'us': round(ms * 1e3, 2)})


def c3(n=5546):
    c, h, w = 4, 350, 525
    tags = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]
    srcs, gt = synth(n, c, h, w, 4, 3, 0.5)
    thr = ops.threshold_grid(100)
    out = torch.empty((n, c, 2, 101), dtype=torch.int32, device=DEV)
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=tags, out=out))
    report('C3 cloud 4-flip TTA %dx4x350x525 fp16: seg_hist_guarded_rows' % n, ms,
           n * c * (h * w * (4 * 2 + 1) + 808), n * c)


def c3b(n=1024):
    """C3 with the per-pixel bin planes as second output (what the pipeline
    runs when the masks are to be encoded afterwards)."""
    c, h, w = 4, 350, 525
    tags = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]
    srcs, gt = synth(n, c, h, w, 4, 3, 0.5)
    thr = ops.threshold_grid(100)
    out = torch.empty((n, c, 2, 101), dtype=torch.int32, device=DEV)
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=tags, bins=True))
    report('C3 + bin planes %dx4x350x525 fp16: seg_hist_guarded_rows<bins>'
           % n, ms, n * c * (h * w * (4 * 2 + 1 + 1) + 808), n * c)


def c5(n=3750):
    c, h, w = 4, 256, 1600
    srcs, gt = synth(n, c, h, w, 5, 5, 0.86)
    thr = ops.threshold_grid(100)
    out = torch.empty((n, c, 2, 101), dtype=torch.int32, device=DEV)
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, out=out))
    report('C5 ensemble 5 models %dx4x256x1600 fp16: seg_hist_guarded_pf' % n,
           ms, n * c * (h * w * (5 * 2 + 1) + 808), n * c)


def c2g(n=1024):
    c, h, w = 4, 256, 1600
    srcs, gt = synth(n, c, h, w, 1, 2, 0.86)
    thr = ops.threshold_grid(100)
    out = torch.empty((n, c, 2, 101), dtype=torch.int32, device=DEV)
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, out=out))
    report('C2 fp16 without the per-code table (seg_hist_logit) %d img' % n, ms,
           n * c * (h * w * 3 + 808), n * c)
    x32 = [srcs[0].float()]
    ms = timeit(lambda: ops.seg_hist(x32, gt, thr, out=out))
    report('C2 fp32 logits (seg_hist_logit) %d img' % n, ms,
           n * c * (h * w * 5 + 808), n * c)


def t32(n=512):
    """4-flip TTA of fp32 logits (the reference's own dtype) on the Severstal
    shape: the exact multi-source kernel (no guarded estimate for fp32)."""
    c, h, w = 4, 256, 1600
    tags = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]
    srcs, gt = synth(n, c, h, w, 4, 7, 0.86)
    srcs = [s.float() for s in srcs]
    thr = ops.threshold_grid(100)
    out = torch.empty((n, c, 2, 101), dtype=torch.int32, device=DEV)
    ms = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=tags, out=out))
    report('4-flip TTA fp32 %dx4x256x1600: seg_hist_generic<float>' % n, ms,
           n * c * (h * w * (4 * 4 + 1) + 808), n * c)


def c2r(n=1024):
    """Single-model submission of the Severstal shape: every plane of ONE fp16
    source to run-length pairs (Infer without TTA): rle_kernel<half, CODE, 512>."""
    c, h, w = 4, 256, 1600
    srcs, _ = synth(n, c, h, w, 1, 2, 0.86)
    runs, off, cnt, area = ops.rle_encode(srcs, 0.5, 0)
    n_runs = int(cnt.sum().item())
    ms = timeit(lambda: ops.rle_encode(srcs, 0.5, 0))
    report('Severstal single source -> RLE %dx4x256x1600 fp16: rle_kernel CODE'
           % n, ms, n * c * h * w * 2 + 8 * n_runs, n * c, {'runs': n_runs})


def c4(n=10000):
    c, h, w = 1, 1024, 1024
    srcs, _ = synth(n, c, h, w, 1, 4, 0.78)
    lib_cap = n * h * w // 16
    import ctypes
    from mlcomp_b200 import _lib
    lib = _lib.load()
    runs = torch.empty(lib_cap, dtype=torch.int32, device=DEV)
    off = torch.empty(n, dtype=torch.int64, device=DEV)
    cnt = torch.empty(n, dtype=torch.int32, device=DEV)
    area = torch.empty(n, dtype=torch.int32, device=DEV)
    total = torch.zeros(1, dtype=torch.int64, device=DEV)
    ptrs = (ctypes.c_void_p * 1)(srcs[0].data_ptr())
    xf = (ctypes.c_int32 * 1)(0)
    thr = (ctypes.c_float * 1)(0.5)
    ar = (ctypes.c_int32 * 1)(3500)

    def run():
        _lib.check(lib.mlc_rle_encode(
            ptrs, xf, 1, 0, n, c, h, w, thr, ar, 1,
            ctypes.c_void_p(runs.data_ptr()), lib_cap,
            ctypes.c_void_p(off.data_ptr()), ctypes.c_void_p(cnt.data_ptr()),
            ctypes.c_void_p(area.data_ptr()),
            ctypes.c_void_p(total.data_ptr()),
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    ms = timeit(run)
    n_runs = int(cnt.sum().item())
    report('C4 SIIM %dx1x1024x1024 fp16 thr 0.5 min-area 3500: rle_kernel'
           % n, ms, n * h * w * 2 + 8 * n_runs, n,
           {'runs': n_runs, 'total_elems': int(total.item())})


def am(n=1024):
    """Segmentation arg-max: 5 classes (4 + background), 2 flips."""
    c, h, w = 5, 256, 1600
    srcs, gt = synth(n, c, h, w, 2, 6, 0.5)
    tags = [[], ['hflip']]
    ms = timeit(lambda: ops.seg_argmax(srcs, tags, 'softmax', gt_onehot=gt,
                                       want_counts=True))
    report('arg-max %dx5x256x1600 fp16 x 2 flips, softmax: seg_argmax_items'
           % n, ms, n * c * h * w * (2 * 2 + 1) + n * h * w, n * c)


def c4s(n=512):
    """C4 all the way to Python strings: encode + device formatting + D2H +
    slicing (what Infer.submit needs), against mask2rle's 3.8 ms per plane."""
    import time
    c, h, w = 1, 1024, 1024
    srcs, _ = synth(n, c, h, w, 1, 4, 0.78)
    for _ in range(2):
        runs, off, cnt, _a = ops.rle_encode(srcs, 0.5, 3500)
        strs = ops.rle_strings(runs, off, cnt)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    runs, off, cnt, _a = ops.rle_encode(srcs, 0.5, 3500)
    strs = ops.rle_strings(runs, off, cnt)
    dt = time.perf_counter() - t0
    chars = sum(len(x) for x in strs)
    print(json.dumps({'config': 'C4 to strings: %d planes 1024x1024 -> rle '
                      'pairs -> decimal text on device -> host str' % n,
                      'ms_total': round(dt * 1e3, 2),
                      'ms_per_plane': round(dt * 1e3 / n, 4),
                      'chars': chars, 'runs': int(cnt.sum().item())}),
          flush=True)


if __name__ == '__main__':
    which = sys.argv[1:] or ['c1', 'c2g', 'c3', 'c4', 'c5']
    for k in which:
        name, _, count = k.partition(':')
        if count:
            globals()[name](int(count))
        else:
            globals()[name]()
        torch.cuda.empty_cache()
