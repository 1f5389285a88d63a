#!/usr/bin/env python
"""C5 (Equation ensemble: 5 models' logits averaged, 256x1600x4, images
sharded over the ranks, one exchange of the packed score tables) at N GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \\
        --master-addr 127.0.0.1 --master-port 29544 tools/bench_c5_multi.py [images_per_rank]

Each rank owns its shard in HBM (synthetic, SURVEY.md 8(d): members = shared
logits + their own noise per pixel); a step = SegSearch.add over the shard + result() (the all_gather
of 16.5 KB per rank + the fold kernel).  Prints one JSON line on rank 0; time = max over ranks.
"""
import datetime
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from mlcomp_b200 import ops, scoring  # noqa: E402


def main():
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev,
                                timeout=datetime.timedelta(seconds=300))
    # BASELINE config 5 at full size: 30 000 images over 8 GPUs = 3 750 each
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 3750
    cfg = dict(bench.CFG)
    c, h, w = cfg['n_cls'], cfg['H'], cfg['W']
    members, gt = bench.synth_tta(cfg, n, 5000 + rank, dev, n_var=5)
    # ensemble members are not flipped: undo the storage flips of synth_tta
    for k in range(5):
        dims = bench.FLIP_DIMS[k % 4]
        if dims:
            for lo in range(0, n, 256):
                members[k][lo:lo + 256] = members[k][lo:lo + 256].flip(*dims)
    thr = ops.threshold_grid(cfg['bins'])
    search = scoring.SegSearch(c, thr, cfg['min_areas'], device=dev)

    def step():
        search.reset()
        search.add(members, gt)
        return search.result()

    for _ in range(3):
        res = step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    k = 10
    e0.record()
    for _ in range(k):
        res = step()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / k], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        px = n * c * h * w
        nbytes = px * (5 * 2 + 1)
        peak = bench.measured_peak()[0]
        print(json.dumps({
            'config': 'C5 ensemble of 5 models, %d img x %d cls x %dx%d fp16 '
                      'per GPU' % (n, c, h, w),
            'n_gpus': world, 'ms_per_step': ms,
            'masks_per_s': n * c * world / (ms * 1e-3),
            'GB/s_per_gpu': nbytes / (ms * 1e-3) / 1e9,
            'frac_of_measured_peak_per_gpu': nbytes / (ms * 1e-3) / 1e9 / peak,
            'n_img_scored': res['n_img'],
            'best': res['best']}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
