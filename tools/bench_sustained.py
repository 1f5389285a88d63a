#!/usr/bin/env python
"""Sustained timing of the 4-flip TTA histogram pass (with bin planes):
back-to-back launches for about a second, SM clock sampled meanwhile - under
the 1 kW power cap a compute-heavy kernel does not hold the 1965 MHz a burst of
a few launches sees.  MLC_LIB picks a library variant."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import CFG, TAGS, ClockSampler, synth_tta  # noqa: E402
from mlcomp_b200 import ops  # noqa: E402

DEV = torch.device('cuda', 0)
TAG = os.path.basename(os.environ.get('MLC_LIB', 'default'))

if __name__ == '__main__':
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 150
    srcs, gt = synth_tta(dict(CFG), n, 7, DEV)
    thr = ops.threshold_grid(100)
    for bins in (True, False):
        for _ in range(3):
            ops.seg_hist(srcs, gt, thr, xforms=TAGS, bins=bins)
        torch.cuda.synchronize()
        s = ClockSampler(0)
        s.start()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(reps):
            ops.seg_hist(srcs, gt, thr, xforms=TAGS, bins=bins)
        e1.record()
        torch.cuda.synchronize()
        clk = s.stop()
        ms = e0.elapsed_time(e1) / reps
        px = n * 4 * 256 * 1600
        gbs = (px * 9 + n * 4 * 808) / (ms * 1e-3) / 1e9
        print(json.dumps({'tag': TAG, 'case': 'tta4_%dimg_%s_sustained' %
                          (n, 'hist+bins' if bins else 'hist'),
                          'ms': round(ms, 4), 'GB/s': round(gbs, 1),
                          'frac_of_measured_peak': round(gbs / 6553.3, 3),
                          'seconds': round(ms * reps / 1e3, 2),
                          'clocks': clk}), flush=True)
