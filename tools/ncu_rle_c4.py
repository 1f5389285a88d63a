"""One RLE launch on C4-shaped planes (1024x1024 fp16, one class), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sys.argv = sys.argv[:1]
import importlib.util
spec = importlib.util.spec_from_file_location(
    'bc', os.path.join(os.path.dirname(os.path.abspath(__file__)), 'bench_configs.py'))
bc = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bc)
from mlcomp_b200 import ops
n = 512
logits, _ = bc.synth(n, 1, 1024, 1024, 1, 4, 0.78)
cap = n * 1024 * 1024 // 16
for _ in range(3):
    ops.rle_encode(logits, [0.5], [3500], capacity=cap)
torch.cuda.synchronize()
