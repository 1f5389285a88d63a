#!/usr/bin/env python
"""Timings of the auxiliary kernels (seg_prob, rle_decode, rle_format,
plane_dice, mask composites) on Severstal-shaped planes."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from mlcomp_b200 import ops
dev = torch.device('cuda', 0)
PEAK = bench.measured_peak()[0]
def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(iters)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))
def rep(name, ms, nbytes):
    print(json.dumps({'kernel': name, 'ms': round(ms, 4), 'GB/s': round(nbytes / ms / 1e6, 1),
                      'frac_of_measured_peak': round(nbytes / ms / 1e6 / PEAK, 3)}))
cfg = dict(bench.C2)
n = 512
logits, gt = bench.synth(cfg, n, 7, dev)
px = logits.numel()
rep('seg_prob fp16 -> fp32', timeit(lambda: ops.seg_prob([logits])), px * 6)
thr = ops.threshold_grid(100)
hist = ops.seg_hist([logits], gt, thr, lut=ops.seg_lut(thr))
rep('plane_dice', timeit(lambda: ops.plane_dice(hist, [50] * 4, [0] * 4)), hist.numel() * 4)
runs, off, cnt, _ = ops.rle_encode([gt], 0.5, 0, activation=None)
rep('rle_encode uint8 masks (gt)', timeit(lambda: ops.rle_encode([gt], 0.5, 0, activation=None, capacity=runs.numel() + 16)), px)
rep('rle_decode -> uint8 planes', timeit(lambda: ops.rle_decode(runs, off.reshape(-1), cnt.reshape(-1), 256, 1600)), px + runs.numel() * 4)
rep('rle_strings (format + D2H + slicing), %d runs' % (runs.numel() // 2), timeit(lambda: ops.rle_strings(runs, off, cnt), iters=3, warm=1), runs.numel() * 4)
m = gt[:64]
rep('mask_composite 64 img', timeit(lambda: ops.mask_composite(m)), m.numel() + m.numel() // 4 * 3)
img = torch.zeros((64, 256, 1600, 3), dtype=torch.uint8, device=dev)
rep('mask_outline 64 img', timeit(lambda: ops.mask_outline(img, m)), m.numel() + img.numel())
