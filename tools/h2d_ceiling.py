#!/usr/bin/env python
"""Host->device copy ceiling of this box: plain cudaMemcpyAsync from pinned
memory on 1, 2, 4, 8 GPUs at the same time (one stream and one pinned buffer
per GPU, chunks of the size the executor stages).  Writes
profiles/h2d_ceiling.json: {"N": aggregate GB/s over the N GPUs, ...} - the
denominator of bench.py's e2e.frac_of_h2d_ceiling."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def measure(n_gpu, chunk_mb=128, chunks=8, reps=6):
    bufs = []
    for d in range(n_gpu):
        torch.cuda.set_device(d)
        h = torch.empty(chunks * chunk_mb << 20, dtype=torch.uint8,
                        pin_memory=True)
        h.fill_(d + 1)
        g = torch.empty(chunk_mb << 20, dtype=torch.uint8, device=f'cuda:{d}')
        bufs.append((h, g, torch.cuda.Stream(device=d)))
    per = chunk_mb << 20

    def run():
        for d, (h, g, s) in enumerate(bufs):
            with torch.cuda.stream(s):
                for c in range(chunks):
                    g.copy_(h[c * per:(c + 1) * per], non_blocking=True)
        for _, _, s in bufs:
            s.synchronize()
    run()
    best = 0.0
    for _ in range(reps):
        t0 = time.perf_counter()
        run()
        dt = time.perf_counter() - t0
        best = max(best, n_gpu * chunks * per / dt / 1e9)
    return best


if __name__ == '__main__':
    n_avail = torch.cuda.device_count()
    out = {}
    for n in (1, 2, 4, 8):
        if n <= n_avail:
            out[str(n)] = round(measure(n), 2)
    out['note'] = ('aggregate GB/s, pinned host memory, cudaMemcpyAsync in '
                   '128 MiB chunks, best of 6; %d GPUs visible' % n_avail)
    print(json.dumps(out))
    if '--write' in sys.argv:
        path = os.path.join(ROOT, 'gpurun_out', 'h2d_ceiling.json')
        os.makedirs(os.path.dirname(path), exist_ok=True)
        json.dump(out, open(path, 'w'), indent=1)
