#!/usr/bin/env python
"""Headline benchmark: validated masks/sec of the scoring hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json `metric` / north_star target, on configs[1]'s shapes):
Severstal-shaped validation set, 6,666 images x 4 classes x 256x1600, every
image predicted under 4 flip variants (id, h, v, hv; fp16 logits, stored as the
model emitted them, i.e. flipped) + uint8 one-hot ground truth, PER GPU (weak
scaling).  One "step" is the whole pipeline over all planes:

  1. fused un-flip + sigmoid + mean + label-conditioned threshold histogram
     of every plane, all 100 thresholds in one read (also keeps every pixel's
     bin index, 1 B/px)                                    [seg_hist_guarded]
  2. histogram scan -> dice for the 100 x 5 (threshold x min-area) grid, summed
     over images; N>1: one all-gather of the packed partials      [dice_search]
  3. D2H of the score table, arg-max per class on the host
  4. mask area of every plane at the chosen threshold          [plane_area]
  5. RLE of every mask at the chosen (threshold, min-area); planes whose mask
     is empty / rejected are known from 4 and not read            [rle_kernel]
  6. D2H of the run count

A "mask" is one (image, class) plane.  One JSON line on stdout (rank 0).
`--impl reference` times the reference's CPU path (oracle/_ref = the
reference's own dice.py / rle.py / tensors.py when present, else the oracle
port) on all host cores.  The former headline (C2: one source, no RLE) stays
as the side key `c2_single_source`.
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(n_img=6666, n_cls=4, H=256, W=1600, bins=100,
           min_areas=(0, 256, 512, 1024, 2048), empty=0.86)
TAGS = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]
FLIP_DIMS = [(), (3,), (2,), (2, 3)]
METRIC = 'validated_masks_per_sec'
UNIT = 'masks/s'
SIGMOID_NOTE = ('bit-equal to torch CUDA sigmoid (the reference\'s device '
                'path, utils/torch.py:15,24-25); CPU oracle (torch CPU '
                'sigmoid) differs by <= 4 ulp per source')


def workload_name(cfg, n_img=None):
    return ('north_star pipeline on C2 shapes: 4-flip TTA (id,h,v,hv) + 100 '
            'thr x %d min-area dice search + RLE of every mask at the chosen '
            'point; %d img x %d cls x %dx%d fp16 x 4 variants + u8 gt per GPU'
            % (len(cfg['min_areas']), n_img or cfg['n_img'], cfg['n_cls'],
               cfg['H'], cfg['W']))


def config_dict(cfg, n_img, e2e_images, e2e_chunk):
    """The `config` object - identical in both arms (the reference arm's
    bounded sample is described under `cpu_baseline.sample`)."""
    planes = n_img * cfg['n_cls']
    px = cfg['H'] * cfg['W']
    return {
        'workload': workload_name(cfg, n_img),
        'planes_per_gpu': planes,
        'arithmetic': 'fp16 logits, fp32 sigmoid / mean / compares, '
                      'integer counts, fp64 dice',
        'cache': 'inputs %.1f GB per GPU per step, far larger than the '
                 '126 MB L2; no flush needed' % (planes * px * 9 / 1e9),
        'e2e': '%d images x 4 variants per GPU per step from pinned host '
               'memory in parts of %d images, copies double-buffered against '
               'the kernels; scores, run pairs and RLE text read back' %
               (min(n_img, e2e_images), e2e_chunk),
    }


def hist_bytes_per_plane(cfg, n_src, elem=2):
    """logits once + 1 B/px ground truth + histogram out."""
    px = cfg['H'] * cfg['W']
    return n_src * px * elem + px + 2 * (cfg['bins'] + 1) * 4


# ---------------------------------------------------------------------------
# synthetic data: mostly-empty planes, a few ellipses,
# logits = +-4 around the mask + N(0, 1.5); TTA variant = the shared logits +
# its own N(0, 0.5) per pixel, stored flipped (what the model emitted)
# ---------------------------------------------------------------------------
def synth_tta(cfg, n_img, seed, device, chunk=32, n_var=4):
    import torch
    c, h, w = cfg['n_cls'], cfg['H'], cfg['W']
    g = torch.Generator(device=device).manual_seed(seed)
    srcs = [torch.empty((n_img, c, h, w), dtype=torch.float16, device=device)
            for _ in range(n_var)]
    gt = torch.empty((n_img, c, h, w), dtype=torch.uint8, device=device)
    yy = torch.arange(h, device=device, dtype=torch.float32)[None, :, None]
    xx = torch.arange(w, device=device, dtype=torch.float32)[None, None, :]
    for lo in range(0, n_img, chunk):
        hi = min(n_img, lo + chunk)
        p = (hi - lo) * c
        empty = torch.rand(p, generator=g, device=device) < cfg['empty']
        mask = torch.zeros((p, h, w), dtype=torch.bool, device=device)
        for _ in range(2):
            cy = torch.rand(p, generator=g, device=device) * h
            cx = torch.rand(p, generator=g, device=device) * w
            frac = 0.001 + 0.05 * torch.rand(p, generator=g, device=device)
            aspect = 0.5 + 2.5 * torch.rand(p, generator=g, device=device)
            ry = torch.sqrt(frac * h * w / 3.14159 / aspect)
            rx = ry * aspect
            mask |= (((yy - cy[:, None, None]) / ry[:, None, None]) ** 2 +
                     ((xx - cx[:, None, None]) / rx[:, None, None]) ** 2) <= 1
        mask &= ~empty[:, None, None]
        lg = torch.randn((p, h, w), generator=g, device=device,
                         dtype=torch.float32) * 1.5
        lg += mask.float() * 8 - 4
        lg = lg.view(hi - lo, c, h, w)
        for k in range(n_var):
            v = lg + 0.5 * torch.randn(lg.shape, generator=g, device=device,
                                       dtype=torch.float32)
            dims = FLIP_DIMS[k % 4]
            srcs[k][lo:hi] = (v.flip(*dims) if dims else v).to(torch.float16)
            del v
        gt[lo:hi] = mask.to(torch.uint8).view(hi - lo, c, h, w)
        del mask, lg
    return srcs, gt


# ---------------------------------------------------------------------------
# CPU leg: what the reference's Valid + Infer executors did per plane
# ---------------------------------------------------------------------------
_REF = None


def _cpu_init():
    import torch
    torch.set_num_threads(1)        # worker/tasks.py:122-126 sets OMP/MKL = 1
    global _REF
    from oracle import make_ref
    _REF = make_ref.load()


def _cpu_plane(args):
    """One plane: sigmoid per variant (utils/torch.py:15), TtaWrap.inverse
    (contrib/transform/tta.py:20-31 via flip, contrib/torch/tensors.py:4-8),
    mean, dice_numpy at every (t, a) of the grid
    (contrib/metrics/dice.py:4-23), mask2rle at one (t, a)
    (contrib/transform/rle.py:4-13).  The three leaf functions are the
    reference's own files when oracle/_ref exists."""
    import torch
    import oracle
    planes, gt, tags, bins, areas, t_rle, a_rle = args
    ref = _REF
    acc = None
    for x, tg in zip(planes, tags):
        p = torch.sigmoid(torch.from_numpy(x).float()[None, None])
        if ref is not None:
            for t in tg:
                p = ref.flip(p, 3 if t == 'hflip' else 2)
        else:
            p = oracle.tta_inverse(p, tg)
        p = p.numpy()[0, 0]
        acc = p.copy() if acc is None else acc + p
    prob = acc / np.float32(len(planes))
    thr = oracle.threshold_grid(bins)
    g = gt.astype(np.float32)
    table = np.zeros((len(thr), len(areas)))
    dice = ref.dice_numpy if ref is not None else oracle.plane_dice
    for k, t in enumerate(thr):
        for m, a in enumerate(areas):
            table[k, m] = dice(g, prob, threshold=np.float32(t),
                               min_area=int(a))
    mask = (prob >= np.float32(t_rle)).astype(np.uint8)
    if a_rle and mask.sum() < a_rle:
        mask[:] = 0
    rle = ref.mask2rle(mask) if ref is not None else oracle.rle_encode(mask)
    return table, rle, prob


def cpu_run(pool, srcs_np, gt_np, tags, cfg, t_c, a_c):
    """srcs_np: list of [planes, H, W] fp16 (stored flipped), gt_np
    [planes, H, W]; class of plane i = i % n_cls."""
    c = cfg['n_cls']
    tasks = [([s[i] for s in srcs_np], gt_np[i], tags, cfg['bins'],
              list(cfg['min_areas']), t_c[i % c], a_c[i % c])
             for i in range(len(gt_np))]
    t0 = time.perf_counter()
    out = pool.map(_cpu_plane, tasks, chunksize=1)
    dt = time.perf_counter() - t0
    return (len(tasks) / dt, dt, np.stack([o[0] for o in out]),
            [o[1] for o in out], np.stack([o[2] for o in out]))


def cpu_kind():
    from oracle import make_ref
    return 'reference' if make_ref.load() is not None else 'port'


def cpu_model():
    try:
        for line in open('/proc/cpuinfo'):
            if line.startswith('model name'):
                return line.split(':', 1)[1].strip()
    except OSError:
        pass
    return 'unknown'


def make_pool():
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    ctx = mp.get_context('fork')    # forked BEFORE CUDA is initialised
    return ctx.Pool(cores, initializer=_cpu_init), cores


# ---------------------------------------------------------------------------
# clocks during the timed region (NVML polling thread)
# ---------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: 'hw_slowdown', 0x40: 'hw_thermal_slowdown',
               0x20: 'sw_thermal_slowdown', 0x4: 'sw_power_cap',
               0x80: 'hw_power_brake_slowdown'}

    def __init__(self, device_index):
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            if not uuid.startswith('GPU-'):
                uuid = 'GPU-' + uuid
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.nv = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(
                self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # pragma: no cover
            self.nv = None
            self.err = repr(e)

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append(mhz)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': [],
                    'samples': 0}
        return {'sm_mhz': float(np.median(self.samples)),
                'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(self.samples)}


def measured_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        return float(json.load(open(path))['hbm_gbs']), 'measured'
    except Exception:
        return 6650.0, 'fallback'


def committed(name, key):
    try:
        d = json.load(open(os.path.join(ROOT, 'profiles', name)))
        return d.get(key)
    except Exception:
        return None


# ---------------------------------------------------------------------------
def run_reference(args, cfg):
    """The reference's CPU implementation of the path, all host cores, each
    step a bounded sample of the workload."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    pool, cores = make_pool()
    n_planes = max(16, cores)
    n_img = (n_planes + cfg['n_cls'] - 1) // cfg['n_cls']
    srcs, gt = synth_tta(cfg, n_img, 2000, 'cpu', chunk=4)
    h, w = cfg['H'], cfg['W']
    s_np = [s.numpy().reshape(-1, h, w)[:n_planes] for s in srcs]
    g_np = gt.numpy().reshape(-1, h, w)[:n_planes]
    t_c, a_c = [0.5] * cfg['n_cls'], [1024] * cfg['n_cls']
    for _ in range(args.warmup):
        cpu_run(pool, s_np, g_np, TAGS, cfg, t_c, a_c)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_run(pool, s_np, g_np, TAGS, cfg, t_c, a_c)
    dt = time.perf_counter() - t0
    pool.close()
    value = n_planes * args.steps / dt
    kind = cpu_kind()
    sample = ('%d planes per step (%d images) of the workload, one plane per '
              'task over %d processes; %s' %
              (n_planes, n_img, cores,
               'dice_numpy / mask2rle / flip are the reference\'s own files '
               '(oracle/_ref), sigmoid + mean restated' if kind == 'reference'
               else 'oracle port'))
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT,
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * dt / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16',
        'data': 'synthetic',
        'config': config_dict(cfg, args.images or cfg['n_img'],
                              args.e2e_images, args.e2e_chunk),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores,
                         'kind': kind, 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0,
                'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }))


# ---------------------------------------------------------------------------
# parity of this very run against the CPU path
# ---------------------------------------------------------------------------
def parity_check(ops, cfg, srcs, gt, tags, thr, hist, bins, best, rle_out,
                 cpu_table, cpu_rle, cpu_prob, n_planes):
    """Compare the first ``n_planes`` planes of the step's outputs with the
    CPU path's.  Integer outputs must agree exactly, except where a pixel's
    CPU and GPU probabilities fall on different sides of a threshold (the two
    sigmoids differ by <= 4 ulp): those pixels are found and listed, and
    every disagreeing (plane, threshold, area) dice cell / RLE string must be
    explained by one of them."""
    import torch
    c, h, w = cfg['n_cls'], cfg['H'], cfg['W']
    n_img = n_planes // c
    areas = list(cfg['min_areas'])
    nb = len(thr) + 1
    sub = [s[:n_img] for s in srcs]
    # (1) the device's exact probabilities: bins and histograms must be THEIR
    #     binning, bit for bit
    p_gpu = ops.seg_prob(sub, xforms=tags).cpu().numpy().reshape(-1, h, w)
    b_gpu = bins[:n_img].cpu().numpy().reshape(-1, h * w).astype(np.int64)
    h_gpu = hist[:n_img].cpu().numpy().reshape(-1, 2, nb).astype(np.int64)
    g_np = gt[:n_img].cpu().numpy().reshape(-1, h * w)
    exact = True
    strad = []
    n_strad = 0
    plane_strad = {}
    for i in range(n_planes):
        cg = np.searchsorted(thr, p_gpu[i].ravel(), side='right')
        exact &= bool(np.array_equal(cg, b_gpu[i]))
        want = np.bincount((g_np[i] != 0) * nb + cg,
                           minlength=2 * nb).reshape(2, nb)
        exact &= bool(np.array_equal(want, h_gpu[i]))
        cc = np.searchsorted(thr, cpu_prob[i].ravel(), side='right')
        bad = np.nonzero(cg != cc)[0]
        if len(bad):
            n_strad += len(bad)
            plane_strad[i] = [(int(min(cg[j], cc[j])), int(max(cg[j], cc[j])))
                              for j in bad]
            for j in bad[:4]:
                if len(strad) < 16:
                    strad.append({'plane': i, 'y': int(j // w),
                                  'x': int(j % w),
                                  'p_gpu': float(p_gpu[i].ravel()[j]),
                                  'p_cpu': float(cpu_prob[i].ravel()[j])})

    def explained(i, k):
        # a straddler of plane i moves between bins lo < hi: it changes the
        # mask at threshold k when lo <= k < hi
        return any(lo <= k < hi for lo, hi in plane_strad.get(i, ()))
    # (2) dice at every (threshold, area) cell of those planes
    t_, p_, i_ = ops.counts_from_hist(hist[:n_img])
    t_ = t_.cpu().numpy().reshape(-1)
    p_ = p_.cpu().numpy().reshape(n_planes, -1)
    i_ = i_.cpu().numpy().reshape(n_planes, -1)
    import oracle
    cells = bad_cells = unexplained = 0
    for i in range(n_planes):
        for m, a in enumerate(areas):
            got = oracle.dice_from_counts(t_[i], p_[i], i_[i], int(a))
            ref = cpu_table[i, :, m]
            diff = np.abs(got - ref) > 1e-5 * np.maximum(np.abs(ref), 1e-12)
            cells += len(ref)
            for k in np.nonzero(diff)[0]:
                bad_cells += 1
                unexplained += 0 if explained(i, int(k)) else 1
    # (3) RLE strings at the chosen point
    runs, off, cnt = rle_out
    got = ops.rle_strings_host(runs, off.reshape(-1)[:n_planes],
                               cnt.reshape(-1)[:n_planes])
    rle_bad = rle_unexplained = 0
    for i in range(n_planes):
        if got[i] != cpu_rle[i]:
            rle_bad += 1
            rle_unexplained += 0 if explained(
                i, best[i % c]['threshold_index']) else 1
        # against the device's own probabilities: always exact
        exact &= got[i] == oracle.rle_submission(
            p_gpu[i], best[i % c]['threshold'], best[i % c]['min_area'])
    ok = exact and unexplained == 0 and rle_unexplained == 0
    return ok, {
        'subset_planes': n_planes,
        'subset_pixels': n_planes * h * w,
        'exact_vs_device_probabilities': bool(exact),
        'straddling_pixels': n_strad,
        'straddlers_listed': strad,
        'dice_cells_checked': cells,
        'dice_cells_differing': bad_cells,
        'unexplained_cells': unexplained,
        'rle_strings_checked': n_planes,
        'rle_strings_differing': rle_bad,
        'unexplained_rle_strings': rle_unexplained,
        'sigmoid': SIGMOID_NOTE,
    }


# ---------------------------------------------------------------------------
def run_ours(args, cfg):
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    pool = cores = None
    if rank == 0 and not args.no_cpu:
        pool, cores = make_pool()          # fork before CUDA exists

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        import datetime
        # fail in minutes, not in NCCL's default 10, if the ranks ever diverge
        dist.init_process_group('nccl', device_id=dev,
                                timeout=datetime.timedelta(seconds=300))
    from mlcomp_b200 import ops, scoring

    n_img, c = args.images or cfg['n_img'], cfg['n_cls']
    h, w = cfg['H'], cfg['W']
    planes = n_img * c
    px = h * w
    srcs, gt = synth_tta(cfg, n_img, 1000 * 2 + rank, dev)
    thr = ops.threshold_grid(cfg['bins'])
    search = scoring.SegSearch(c, thr, cfg['min_areas'], device=dev,
                               record_events=True)
    cap = max(1 << 20, planes * px // 64)
    # one buffer for the bin planes of every step (10.9 GB: allocated per step
    # it alternates between two cached blocks, and whenever the caching
    # allocator gives one back to the driver the step stalls for 50 - 200 ms)
    bins_buf = torch.empty((n_img, c, h, w), dtype=torch.uint8, device=dev)

    def step():
        search.reset()
        hist, bins = search.add(srcs, gt, xforms=TAGS, keep_bins=bins_buf)
        res = search.result()
        k_c = [b['threshold_index'] for b in res['best']]
        a_c = [b['min_area'] for b in res['best']]
        rle = ops.rle_from_bins(bins, k_c, a_c, capacity=cap, hist=hist)
        return res, hist, bins, rle

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    res, hist, bins, rle = step()
    best = res['best']
    n_runs = int(rle[0].numel() // 2)
    # the run buffer of the later steps: twice what this data needs instead
    # of the size-based first guess (0.7 GB per step)
    cap = max(1 << 20, 4 * n_runs)
    n_nonempty = int((rle[2] > 0).sum().item())

    # ---- CPU path on the first planes of the shard + parity (rank 0) -------
    validated = True
    parity = cpu = None
    if rank == 0 and pool is not None:
        n_cpu = max(64, 4 * cores) // c * c
        k_img = min(n_img, n_cpu // c)
        n_cpu = k_img * c
        s_np = [s[:k_img].cpu().numpy().reshape(-1, h, w) for s in srcs]
        g_np = gt[:k_img].cpu().numpy().reshape(-1, h, w)
        rate, dt, table, cpu_rle, cpu_prob = cpu_run(
            pool, s_np, g_np, TAGS, cfg,
            [b['threshold'] for b in best], [b['min_area'] for b in best])
        pool.close()
        _cpu_init()
        t0 = time.perf_counter()
        _cpu_plane(([s[0] for s in s_np], g_np[0], TAGS, cfg['bins'],
                    list(cfg['min_areas']), best[0]['threshold'],
                    best[0]['min_area']))
        single = 1.0 / (time.perf_counter() - t0)
        kind = cpu_kind()
        cpu = {'value': rate, 'unit': UNIT, 'cores': cores, 'kind': kind,
               'single_core_value': single, 'cpu_model': cpu_model(),
               'sample': '%d planes (%d images) of this run\'s data: 4 x '
                         'sigmoid + un-flip + mean, dice_numpy at %d x %d '
                         'points, mask2rle at the chosen point; one plane per '
                         'task over %d processes, %.1f s wall; %s' %
                         (n_cpu, k_img, cfg['bins'], len(cfg['min_areas']),
                          cores, dt,
                          'leaf functions = the reference\'s own files '
                          '(oracle/_ref)' if kind == 'reference'
                          else 'oracle port')}
        ok, parity = parity_check(ops, cfg, srcs, gt, TAGS, thr, hist, bins,
                                  best, rle[:3], table, cpu_rle, cpu_prob,
                                  n_cpu)
        validated &= ok
        del s_np, g_np, table, cpu_prob
    del hist, bins, rle

    # ---- warm-up + timed region ---------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step()
    sampler = ClockSampler(local) if rank == 0 and not os.environ.get('MLC_BENCH_NO_SAMPLER') else None
    search.kernel_events.clear()
    barrier()
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    # no collector pauses inside the timed steps (with several ranks one
    # rank's pause is everybody's: the ranks meet in every step's exchange)
    import gc
    gc.collect()
    gc.disable()
    # no-ops unless run under `ncu --profile-from-start off`
    torch.cuda.profiler.start()
    e0.record()
    marks = [e0]
    for _ in range(args.steps):
        out = step()
        marks.append(torch.cuda.Event(True))
        marks[-1].record()
    e1.record()
    barrier()
    gc.enable()
    torch.cuda.profiler.stop()
    clocks = sampler.stop() if sampler else None
    ms = e0.elapsed_time(e1)
    each_ms = [a.elapsed_time(b) for a, b in zip(marks[:-1], marks[1:])]
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = planes * world * args.steps / (ms * 1e-3)
    kern_ms = [a.elapsed_time(b) for a, b in search.kernel_events]
    kern_avg = float(np.mean(kern_ms))
    # the slowest step / launch of every rank (a one-off stall on one GPU
    # shows here and in the max-over-ranks total, not in the medians)
    worst = torch.tensor([max(each_ms), float(np.median(each_ms)),
                          max(kern_ms), float(np.median(kern_ms))],
                         dtype=torch.float64, device=dev)
    if world > 1:
        allw = [torch.zeros_like(worst) for _ in range(world)]
        dist.all_gather(allw, worst)
        worst = torch.stack(allw)
    else:
        worst = worst[None]
    worst = worst.cpu().numpy().round(3).tolist()
    same_best = [b['threshold_index'] for b in out[0]['best']] == \
        [b['threshold_index'] for b in best]
    validated &= same_best
    del out

    side = {}
    # ---- strong scaling: the SAME 6,666-image set split over the ranks -----
    if world > 1:
        lo, hi = scoring.shard_range(n_img, rank, world)
        m = hi - lo
        part = [s_[:m] for s_ in srcs]
        gpart = gt[:m]

        def strong_step():
            search.reset()
            h_, b_ = search.add(part, gpart, xforms=TAGS, keep_bins=bins_buf[:m])
            r_ = search.result()
            return ops.rle_from_bins(
                b_, [b['threshold_index'] for b in r_['best']],
                [b['min_area'] for b in r_['best']], capacity=cap, hist=h_)
        for _ in range(3):
            strong_step()
        barrier()
        e0.record()
        for _ in range(args.steps):
            strong_step()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64,
                         device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_s = float(t.item()) / args.steps
        side['strong_scaling'] = {
            'workload': 'the %d-image set split over %d ranks '
                        '(scoring.shard_range), whole pipeline + exchange' %
                        (n_img, world),
            'images_per_rank': m, 'ms_per_step': ms_s,
            'value': planes / (ms_s * 1e-3), 'unit': UNIT,
            'speedup_vs_one_gpu_step': (ms / args.steps) / ms_s}
    # ---- side: the pass without bins, RLE re-reading the sources -----------
    if not args.no_side:
        def step_reread():
            search.reset()
            hist = search.add(srcs, gt, xforms=TAGS)
            r = search.result()
            k_c = [b['threshold_index'] for b in r['best']]
            return ops.rle_encode(
                srcs, [b['threshold'] for b in r['best']],
                [b['min_area'] for b in r['best']], xforms=TAGS, capacity=cap,
                known_area=ops.plane_area(hist, k_c))
        for _ in range(2):
            rr = step_reread()
        barrier()
        e0.record()
        for _ in range(max(3, args.steps // 4)):
            rr = step_reread()
        e1.record()
        barrier()
        ms_rr = e0.elapsed_time(e1) / max(3, args.steps // 4)
        side['pipeline_rereading_sources'] = {
            'ms_per_step': ms_rr, 'value': planes / (ms_rr * 1e-3),
            'same_runs': int(rr[0].numel() // 2) == n_runs}
        validated &= int(rr[0].numel() // 2) == n_runs
        del rr

    # ---- side: former headline, C2 = one untransformed source, no RLE ------
    if not args.no_side:
        c2 = scoring.SegSearch(c, thr, cfg['min_areas'], device=dev,
                               record_events=True)

        def c2_step():
            c2.reset()
            c2.add(srcs[0], gt)
            return c2.result()
        for _ in range(3):
            c2_step()
        c2.kernel_events.clear()
        barrier()
        e0.record()
        for _ in range(args.steps):
            c2_step()
        e1.record()
        barrier()
        ms_c2 = e0.elapsed_time(e1) / args.steps
        k2 = float(np.mean([a.elapsed_time(b) for a, b in c2.kernel_events]))
        by2 = hist_bytes_per_plane(cfg, 1) * planes
        side['c2_single_source'] = {
            'workload': 'C2: %d img x %d cls x %dx%d fp16, one source, '
                        'histogram + search (no TTA, no RLE)' %
                        (n_img, c, h, w),
            'value': planes / (ms_c2 * 1e-3), 'ms_per_step': ms_c2,
            'kernel': 'seg_hist_cell', 'avg_launch_ms': k2,
            'achieved_gbs': by2 / (k2 * 1e-3) / 1e9,
            'frac': by2 / (k2 * 1e-3) / 1e9 / measured_peak()[0]}
        del c2

    # ---- end to end through the executor plugin -----------------------------
    e2e = e2e_through_executor(args, cfg, srcs, gt, dev, world, barrier)
    validated &= e2e.pop('ok')
    e2e.pop('how', None)

    # ---- side: the other BASELINE configs, kernel timings (N = 1 only) ------
    if world == 1 and not args.no_side:
        del srcs, gt, bins_buf
        torch.cuda.empty_cache()
        try:
            sys.path.insert(0, os.path.join(ROOT, 'tools'))
            import bench_configs as bc
            bc.DEV, bc.QUIET = dev, True
            bc.RESULTS.clear()
            for name, count in (('c1', None), ('c3', 1024), ('c4', 2048),
                                ('c5', 512), ('am', 256)):
                getattr(bc, name)(*([count] if count else []))
                torch.cuda.empty_cache()
            side['other_configs'] = {
                'note': 'kernel timings (median of 10 launches, CUDA events) '
                        'of the other BASELINE.json configs at reduced image '
                        'counts; full sizes: tools/bench_configs.py, '
                        'profiles/r02_kernels.jsonl',
                'lines': list(bc.RESULTS)}
        except Exception as e:  # a side key must not cost the headline
            side['other_configs'] = {'error': repr(e)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_kind = measured_peak()
    bytes_launch = hist_bytes_per_plane(cfg, 4) * planes
    achieved = bytes_launch / (kern_avg * 1e-3) / 1e9
    step_bytes = bytes_launch + 8 * n_runs
    step_ms = ms / args.steps
    out = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world,
        'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': step_ms, 'higher_is_better': True,
        'ms_each_step_rank0': {'min': min(each_ms), 'max': max(each_ms),
                               'median': float(np.median(each_ms))},
        'ms_per_rank': {'columns': ['slowest step', 'median step',
                                    'slowest launch', 'median launch'],
                        'rows': worst},
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f16',
        'data': 'synthetic',
        'config': config_dict(cfg, n_img, args.e2e_images, args.e2e_chunk),
        'validated': bool(validated),
        'parity': dict(parity or {}, best=best,
                       same_best_every_step=same_best),
        'roofline': {
            'bound': 'hbm', 'kernel': 'seg_hist_guarded<rows,bins>',
            'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
            'frac': achieved / peak, 'peak_kind': peak_kind,
            'frac_of_8000_nominal': achieved / 8000.0,
            'algorithmic_bytes_per_launch': bytes_launch,
            'algorithmic_bytes_per_plane': hist_bytes_per_plane(cfg, 4),
            'avg_launch_ms': kern_avg,
            'traffic': committed('traffic.json', 'seg_hist_guarded_pipeline'),
            'whole_step': {
                'algorithmic_bytes': step_bytes,
                'achieved': step_bytes / (step_ms * 1e-3) / 1e9,
                'frac': step_bytes / (step_ms * 1e-3) / 1e9 / peak,
                'note': '(4x2 + 1) B/px + 808 B per plane + 8 B per run, over '
                        'the whole step incl. search, D2H and RLE; the bin '
                        'plane (1 B/px written, read for the %d non-empty '
                        'masks) is extra traffic, not algorithmic bytes' %
                        n_nonempty},
        },
        'pipeline': {'run_pairs': n_runs, 'nonempty_masks': n_nonempty,
                     'planes': planes},
        'cpu_baseline': cpu,
        'e2e': e2e,
        'gpu_launches': 5 * args.steps,
        'clocks': clocks,
    }
    out.update(side)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def e2e_through_executor(args, cfg, srcs, gt, dev, world, barrier):
    """The same pipeline through the plugin API a DAG YAML reaches:
    Executor.from_config(...) -> ValidSegment -> executor(task=, ...).
    Inputs start in pinned host memory; every step copies them to the device
    (double-buffered, overlapped with the kernels) and reads the scores and
    the RLE strings back."""
    import torch
    import torch.distributed as dist
    from mlcomp_b200.worker.executors import Executor
    from mlcomp_b200.worker import store
    c, h, w = cfg['n_cls'], cfg['H'], cfg['W']
    e_img = min(srcs[0].shape[0], args.e2e_images)
    host = [torch.empty((e_img, c, h, w), dtype=torch.float16,
                        pin_memory=True).copy_(s[:e_img]) for s in srcs]
    h_gt = torch.empty((e_img, c, h, w), dtype=torch.uint8,
                       pin_memory=True).copy_(gt[:e_img])
    torch.cuda.synchronize()
    config = {'executors': {'valid': {
        'type': 'valid_segment', 'logits': host, 'tta': TAGS,
        'targets': h_gt, 'bins': cfg['bins'],
        'min_areas': list(cfg['min_areas']), 'part_size': args.e2e_chunk,
        'prepare_submit': True}}}

    def e2e_step():
        ex = Executor.from_config(executor='valid', config=config,
                                  additional_info={}, session=None,
                                  logger=None)
        out = ex(task=store.Task(name='bench'), task_provider=None,
                 dag=store.Dag(name='bench'))
        return ex, out

    for _ in range(2):
        ex, r = e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(args.e2e_steps):
        ex, r = e2e_step()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sec = float(t.item()) * 1e-3
    value = e_img * c * world * args.e2e_steps / sec
    h2d = e_img * c * h * w * (4 * 2 + 1)
    d2h = int(ex.d2h_bytes)
    # the executor's answer must be the device-resident pipeline's answer on
    # the same images
    from mlcomp_b200 import ops, scoring
    s2 = scoring.SegSearch(c, ops.threshold_grid(cfg['bins']),
                           cfg['min_areas'], device=dev)
    s2.reset()
    s2.add([s[:e_img] for s in srcs], gt[:e_img], xforms=TAGS)
    want = s2.result()
    ok = [b['threshold_index'] for b in want['best']] == \
        [b['threshold_index'] for b in ex.best] and \
        abs(want['dice'] - r['res']) < 1e-9
    # the ceiling of THIS box: every rank copies one of its pinned source
    # arrays to the device at the same time, plain cudaMemcpyAsync (the
    # committed figure of tools/h2d_ceiling.py is from another box: boxes
    # differ by 30 % in what N GPUs together get from the host)
    dst = torch.empty_like(host[0], device=dev)
    dst.copy_(host[0], non_blocking=True)
    barrier()
    c0, c1 = torch.cuda.Event(True), torch.cuda.Event(True)
    c0.record()
    for _ in range(3):
        dst.copy_(host[0], non_blocking=True)
    c1.record()
    barrier()
    rate = torch.tensor([3 * host[0].numel() * 2 / (c0.elapsed_time(c1) * 1e-3)
                         / 1e9], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(rate, op=dist.ReduceOp.SUM)
    ceiling = float(rate.item())
    del dst
    gbs = h2d * world * args.e2e_steps / sec / 1e9
    return {
        'value': value, 'unit': UNIT, 'h2d_bytes_per_step': h2d,
        'd2h_bytes_per_step': d2h, 'ok': bool(ok),
        'h2d_gbs_all_gpus': gbs,
        'h2d_ceiling_gbs_all_gpus': ceiling,
        'h2d_ceiling_measured': 'in this run: all %d ranks copying 1.7 GB of '
                                'pinned memory at once, 3 times' % world,
        'h2d_ceiling_committed_other_box':
            committed('h2d_ceiling.json', str(world)),
        'frac_of_h2d_ceiling': (gbs / ceiling) if ceiling else None,
        'through': 'Executor.from_config -> ValidSegment(prepare_submit) '
                   '-> executor(task=, task_provider=, dag=)',
        'how': '%d images x 4 variants per GPU per step from pinned host '
               'memory in parts of %d images, copies double-buffered against '
               'the kernels; scores, run pairs and RLE text read back' %
               (e_img, args.e2e_chunk),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--images', type=int, default=0,
                    help='images per GPU (default: the full set, 6666)')
    ap.add_argument('--e2e-images', type=int, default=512)
    ap.add_argument('--e2e-chunk', type=int, default=32)
    ap.add_argument('--e2e-steps', type=int, default=5)
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-side', action='store_true',
                    help='skip the side measurements')
    args = ap.parse_args()
    cfg = dict(CFG)
    if args.impl == 'reference':
        run_reference(args, cfg)
    else:
        run_ours(args, cfg)


if __name__ == '__main__':
    main()
