"""One 4-flip TTA histogram launch on Severstal-shaped planes (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mlcomp_b200 import ops
dev = torch.device('cuda', 0)
cfg = dict(bench.C2)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
logits, gt = bench.synth(cfg, n, 7, dev)
g = torch.Generator(device=dev).manual_seed(3)
dims = ((), (3,), (2,), (2, 3))
srcs = [(logits + 0.5 * torch.randn(logits.shape, generator=g, device=dev, dtype=torch.float16)).flip(*d).contiguous()
        if d else logits.clone() for d in dims]
xf = [ops.compose_tta(t) for t in ([], ['hflip'], ['vflip'], ['hflip', 'vflip'])]
thr = ops.threshold_grid(100)
for _ in range(3):
    ops.seg_hist(srcs, gt, thr, xforms=xf)
torch.cuda.synchronize()
