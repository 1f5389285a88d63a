"""One single-source RLE launch on Severstal-shaped planes (for ncu captures)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from mlcomp_b200 import ops
dev = torch.device('cuda', 0)
cfg = dict(bench.C2)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
logits, gt = bench.synth(cfg, n, 7, dev)
cap = n * 4 * 256 * 1600 // 32
for _ in range(3):
    ops.rle_encode([logits], [0.5] * 4, [256] * 4, capacity=cap)
torch.cuda.synchronize()
