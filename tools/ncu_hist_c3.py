"""One 4-flip TTA histogram launch on cloud-shaped planes (350x525), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mlcomp_b200 import ops
dev = torch.device('cuda', 0)
n, c, h, w = 256, 4, 350, 525
g = torch.Generator(device=dev).manual_seed(3)
base = (torch.randn((n, c, h, w), generator=g, device=dev) * 1.5 - 4).half()
gt = (torch.rand((n, c, h, w), generator=g, device=dev) < 0.1).to(torch.uint8)
dims = ((), (3,), (2,), (2, 3))
srcs = [(base + 0.5 * torch.randn(base.shape, generator=g, device=dev, dtype=torch.float16)).flip(*d).contiguous()
        if d else base.clone() for d in dims]
xf = [ops.compose_tta(t) for t in ([], ['hflip'], ['vflip'], ['hflip', 'vflip'])]
thr = ops.threshold_grid(100)
for _ in range(3):
    ops.seg_hist(srcs, gt, thr, xforms=xf)
torch.cuda.synchronize()
