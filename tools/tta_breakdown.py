import sys, os, json
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import bench
from mlcomp_b200 import ops, scoring
dev = torch.device('cuda', 0)
cfg = dict(bench.C2)
n = 512
logits, gt = bench.synth(cfg, n, 7, dev)
thr = ops.threshold_grid(100)
xf = [ops.compose_tta(t) for t in ([], ['hflip'], ['vflip'], ['hflip', 'vflip'])]
g=torch.Generator(device=dev).manual_seed(3)
srcs=[(logits+0.5*torch.randn(logits.shape,generator=g,device=dev,dtype=torch.float16)).flip(*d).contiguous() if d else logits.clone() for d in ((),(3,),(2,),(2,3))]
def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(iters)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))
px = n * 4 * 256 * 1600
lut = ops.seg_lut(thr)
r = {}
r['hist_tta4_rows'] = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=xf))
r['hist_4src_flat'] = timeit(lambda: ops.seg_hist(srcs, gt, thr))
r['hist_1src_lut'] = timeit(lambda: ops.seg_hist([logits], gt, thr, lut=lut))
hist = ops.seg_hist(srcs, gt, thr, xforms=xf)
r['dice_search'] = timeit(lambda: ops.dice_search(hist, np.array([[0, 256, 512, 1024, 2048]] * 4)))
cap = px // 32
r['rle_tta4'] = timeit(lambda: ops.rle_encode(srcs, [0.5] * 4, [256] * 4, xforms=xf, capacity=cap))
r['rle_4src_flat'] = timeit(lambda: ops.rle_encode(srcs, [0.5] * 4, [256] * 4, capacity=cap))
r['rle_1src'] = timeit(lambda: ops.rle_encode([logits], [0.5] * 4, [256] * 4, capacity=cap))
for k, v in r.items():
    print(k, round(v, 3), 'ms')
print('GB hist4', px * 9 / 1e9, 'rle4', px * 8 / 1e9, 'hist1', px * 3 / 1e9, 'rle1', px * 2 / 1e9)
same = [logits.flip(*d).contiguous() if d else logits for d in ((), (3,), (2,), (2, 3))]
print('rle_tta4_identical', round(timeit(lambda: ops.rle_encode(same, [0.5] * 4, [256] * 4, xforms=xf, capacity=cap)), 3))
print('rle_hflip_1src', round(timeit(lambda: ops.rle_encode([same[1]], [0.5] * 4, [256] * 4, xforms=[xf[1]], capacity=cap)), 3))
# fp32 logits (the reference's own dtype), 256 images
n32 = 256
s32 = [s[:n32].float() for s in srcs]
px32 = n32 * 4 * 256 * 1600
t = timeit(lambda: ops.seg_hist(s32, gt[:n32], thr, xforms=xf))
print('hist_tta4_rows_fp32', round(t, 3), 'ms', round(px32 * 17 / t / 1e6 / 6550.4, 3), 'of HBM peak')
t = timeit(lambda: ops.seg_hist(s32, gt[:n32], thr))
print('hist_4src_flat_fp32', round(t, 3), 'ms', round(px32 * 17 / t / 1e6 / 6550.4, 3), 'of HBM peak')
t = timeit(lambda: ops.rle_encode(s32, [0.5] * 4, [256] * 4, xforms=xf, capacity=cap))
print('rle_tta4_fp32', round(t, 3), 'ms', round(px32 * 16 / t / 1e6 / 6550.4, 3), 'of HBM peak')
t = timeit(lambda: ops.rle_encode([s32[0]], [0.5] * 4, [256] * 4, capacity=cap))
print('rle_1src_fp32', round(t, 3), 'ms', round(px32 * 4 / t / 1e6 / 6550.4, 3), 'of HBM peak')
# per-pixel bins: cost of writing them in pass 1, and RLE from them
t = timeit(lambda: ops.seg_hist(srcs, gt, thr, xforms=xf, bins=True))
print('hist_tta4_rows_with_bins', round(t, 3), 'ms')
_, bins = ops.seg_hist(srcs, gt, thr, xforms=xf, bins=True)
t = timeit(lambda: ops.rle_from_bins(bins, [50] * 4, [256] * 4, capacity=cap))
print('rle_from_bins', round(t, 3), 'ms')
