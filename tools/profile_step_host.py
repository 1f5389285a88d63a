#!/usr/bin/env python
"""Host-side profile of one headline step (reset + histogram/bins + search +
RLE of the chosen masks): where the ~0.5 ms of a step that is not kernel time
goes.  cProfile over 40 steps on a 512-image set; sync waits show up inside
the calls that block (.cpu(), .item())."""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mlcomp_b200 import ops, scoring  # noqa: E402


def main():
    cfg = bench.CFG
    dev = torch.device('cuda', 0)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    srcs, gt = bench.synth_tta(cfg, n, 2000, dev)
    thr = ops.threshold_grid(cfg['bins'])
    search = scoring.SegSearch(cfg['n_cls'], thr, cfg['min_areas'], device=dev)
    bins_buf = torch.empty((n, cfg['n_cls'], cfg['H'], cfg['W']),
                           dtype=torch.uint8, device=dev)
    cap = 1 << 24

    def step():
        search.reset()
        hist, bins = search.add(srcs, gt, xforms=bench.TAGS, keep_bins=bins_buf)
        res = search.result()
        k_c = [b['threshold_index'] for b in res['best']]
        a_c = [b['min_area'] for b in res['best']]
        return ops.rle_from_bins(bins, k_c, a_c, capacity=cap, hist=hist)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(40):
        step()
    e1.record()
    torch.cuda.synchronize()
    print('ms per step (device time): %.3f' % (e0.elapsed_time(e1) / 40))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(40):
        step()
    torch.cuda.synchronize()
    pr.disable()
    st = pstats.Stats(pr)
    st.sort_stats('cumulative').print_stats(28)


if __name__ == '__main__':
    main()
