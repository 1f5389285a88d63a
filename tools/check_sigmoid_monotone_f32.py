"""Exhaustive check that the device sigmoid (mlc_seg_prob, the arithmetic all
kernels share) is non-decreasing over ALL fp32 inputs (NaNs excluded)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlcomp_b200 import ops
dev = torch.device('cuda', 0)
CH = 1 << 26
bad = 0
prev_last = None
# positive half: bit patterns 0 .. 0x7F800000 ascending = values ascending
def sweep(lo, hi, negative):
    global bad
    last = None
    for b0 in range(lo, hi, CH):
        n = min(CH, hi - b0)
        bits = torch.arange(b0, b0 + n, device=dev, dtype=torch.int64)
        if negative:
            bits = bits + 0x80000000
        x = bits.to(torch.int32 if not negative else torch.int64)
        x = (bits & 0xFFFFFFFF).to(torch.int64)
        x = torch.where(x >= 2**31, x - 2**32, x).to(torch.int32).view(torch.float32)
        p = ops.seg_prob([x.view(1, 1, 1, n)], activation='sigmoid').view(-1)
        d = p[1:] - p[:-1]
        viol = (d > 0) if negative else (d < 0)
        bad += int(viol.sum())
        if last is not None:
            bad += int((p[0] > last) if negative else (p[0] < last))
        last = p[-1].clone()
    return
sweep(0, 0x7F800001, False)      # +0 .. +inf
sweep(0, 0x7F800001, True)       # -0 .. -inf (values descending)
x = torch.tensor([-0.0, 0.0], device=dev).view(1, 1, 1, 2)
p = ops.seg_prob([x], activation='sigmoid').view(-1)
print('violations', bad, 'sigmoid(-0), sigmoid(+0) =', p.tolist())
