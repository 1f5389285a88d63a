"""Dataset-level scoring on top of the kernels: the loops the reference's
(missing) ``Valid`` / ``Infer`` executors ran around ``dice_numpy`` and
``mask2rle``, kept on the device.

``SegSearch``   threshold x min-area dice search over a validation set that is
                fed chunk by chunk (device tensors, or pinned host buffers
                with copies overlapped against compute), optionally sharded
                over ranks and combined with one small all-reduce.
``ClassifyScore`` arg-max, confusion matrix, accuracy (config 1).
"""
import numpy as np
import torch

from . import ops

__all__ = ['SegSearch', 'ClassifyScore', 'HostStager', 'shard_range',
           'combine_ranks', 'combine_packed']


def _xf_code(x):
    return x if isinstance(x, int) else ops.compose_tta(x)


def shard_range(n, rank, world):
    """Contiguous image range of ``rank``: [lo, hi)."""
    return (n * rank) // world, (n * (rank + 1)) // world


def combine_packed(packed, n_f64, group=None):
    """The one exchange step of the sharded path on a packed
    int64 buffer: the first ``n_f64`` entries are fp64 bit patterns (the
    ``dice_sum`` partials), the rest integers (``micro`` counts, image count).

    ONE collective (all_gather of the buffer) and ONE kernel
    (mlc_combine_ranks): integers are summed (exact, order-independent), the
    fp64 partial sums are added in rank order on every rank, so every rank
    (and every rerun at the same world size) holds bit-identical scores.
    NCCL for CUDA tensors; CPU tensors (the gloo tests of the host logic) are
    folded with the same rule by torch.  No-op without an initialised process
    group or at world size 1."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or \
            dist.get_world_size(group) == 1:
        return packed
    world = dist.get_world_size(group)
    n = packed.numel()
    flat = torch.empty(world * n, dtype=torch.int64, device=packed.device)
    dist.all_gather_into_tensor(flat, packed.contiguous().reshape(-1),
                                group=group)
    rows = flat.view(world, n)
    out = torch.empty_like(packed)
    if packed.is_cuda:
        from . import _lib
        lib = _lib.load()
        with torch.cuda.device(packed.device):
            _lib.check(lib.mlc_combine_ranks(
                rows.data_ptr(), world, n_f64, n - n_f64, out.data_ptr(),
                torch.cuda.current_stream().cuda_stream))
        return out
    parts = rows[:, :n_f64].contiguous().view(torch.float64)
    total = parts[0].clone()
    for r in range(1, world):
        total += parts[r]
    out[:n_f64] = total.view(torch.int64)
    out[n_f64:] = rows[:, n_f64:].sum(dim=0)
    return out


def combine_ranks(dice_sum, micro, n_img, group=None):
    """``combine_packed`` for separate tensors: returns the combined
    (dice_sum, micro, n_img)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or \
            dist.get_world_size(group) == 1:
        return dice_sum, micro, n_img
    nd, nm = dice_sum.numel(), micro.numel()
    packed = torch.cat([dice_sum.contiguous().view(torch.int64).reshape(-1),
                        micro.reshape(-1),
                        torch.tensor([n_img], dtype=torch.int64,
                                     device=micro.device)])
    out = combine_packed(packed, nd, group)
    return (out[:nd].view(torch.float64).reshape(dice_sum.shape),
            out[nd:nd + nm].reshape(micro.shape), out[-1])


class HostStager:
    """Host -> device staging of image ranges, copies overlapped with compute.

    ``arrays``: host arrays sharing their first dimension (torch CPU tensors,
    numpy arrays, ``np.load(mmap_mode='r')`` maps).  ``fetch(lo, hi)`` returns
    device views of rows [lo, hi) of every array; ``prefetch(lo, hi)`` starts
    that copy ahead of time on a side stream (``n_buf`` device buffer sets in
    rotation, so the copy of part i+1 runs while part i is being scored);
    ``release()`` marks the most recently fetched buffer set as consumed by the
    work queued so far on the current stream.

    Pinned torch tensors are copied straight from where they lie.  Anything
    else (pageable memory, memory-mapped .npy files) goes through pinned
    bounce buffers filled by a few threads (numpy's copy releases the GIL), so
    the DMA engine is never handed pageable memory.

    Safe across successive passes over the data: a buffer set is only
    overwritten after the copy stream has waited for the event recorded at its
    last ``release()``.
    """

    def __init__(self, arrays, device, part, n_buf=2, threads=8):
        self.arrays = list(arrays)
        self.device = torch.device(device)
        self.part = int(part)
        self.stream = torch.cuda.Stream(device=self.device)
        self.pinned = [torch.is_tensor(a) and a.is_pinned()
                       for a in self.arrays]
        self.bufs = []
        for _ in range(n_buf):
            dev = [torch.empty((self.part,) + tuple(a.shape[1:]),
                               dtype=self._dtype(a), device=self.device)
                   for a in self.arrays]
            host = [None if pin else
                    torch.empty((self.part,) + tuple(a.shape[1:]),
                                dtype=self._dtype(a), pin_memory=True)
                    for a, pin in zip(self.arrays, self.pinned)]
            self.bufs.append({'dev': dev, 'host': host,
                              'copied': torch.cuda.Event(),
                              'consumed': torch.cuda.Event(),
                              'has_copied': False, 'has_consumed': False,
                              'range': None})
        self.next = 0
        self.last = None
        self.pool = None
        if not all(self.pinned):
            from concurrent.futures import ThreadPoolExecutor
            self.pool = ThreadPoolExecutor(max_workers=threads)
            self.threads = threads
        self.h2d_bytes = 0

    @staticmethod
    def _dtype(a):
        if torch.is_tensor(a):
            return a.dtype
        return torch.from_numpy(np.empty(0, dtype=a.dtype)).dtype

    def _bounce(self, dst, a, lo, hi):
        """rows [lo, hi) of a pageable / mapped array -> pinned dst."""
        n = hi - lo
        out = dst[:n].numpy()
        src = a[lo:hi]
        if torch.is_tensor(src):
            src = src.numpy()
        k = max(1, min(self.threads, n))
        cuts = [(n * i) // k for i in range(k + 1)]
        list(self.pool.map(
            lambda i: np.copyto(out[cuts[i]:cuts[i + 1]],
                                src[cuts[i]:cuts[i + 1]]), range(k)))

    def prefetch(self, lo, hi):
        for b in self.bufs:
            if b['range'] == (lo, hi):
                return b
        b = self.bufs[self.next]
        self.next = (self.next + 1) % len(self.bufs)
        n = hi - lo
        if n > self.part:
            raise ValueError('part larger than the staging buffers')
        if any(not p for p in self.pinned) and b['has_copied']:
            b['copied'].synchronize()   # its bounce buffers are free again
        for a, pin, hb in zip(self.arrays, self.pinned, b['host']):
            if not pin:
                self._bounce(hb, a, lo, hi)
        with torch.cuda.stream(self.stream):
            if b['has_consumed']:
                self.stream.wait_event(b['consumed'])
            for a, pin, hb, d in zip(self.arrays, self.pinned, b['host'],
                                     b['dev']):
                d[:n].copy_(a[lo:hi] if pin else hb[:n], non_blocking=True)
                self.h2d_bytes += n * d[0].numel() * d.element_size()
            b['copied'].record(self.stream)
            b['has_copied'] = True
        b['range'] = (lo, hi)
        return b

    def fetch(self, lo, hi):
        b = self.prefetch(lo, hi)
        torch.cuda.current_stream(self.device).wait_event(b['copied'])
        b['range'] = None            # a later pass must copy again
        self.last = b
        return [d[:hi - lo] for d in b['dev']]

    def release(self):
        if self.last is not None:
            self.last['consumed'].record(
                torch.cuda.current_stream(self.device))
            self.last['has_consumed'] = True
            self.last = None


class SegSearch:
    """Accumulates sum_i dice(i, c, t, a) and the global integer counts.

    score[c, t, a] = mean over images of dice_numpy(gt, prob, thr[t], area[a])
    (contrib/metrics/dice.py:4-23, empty_one rule per plane); best point per
    class = first arg-max over (t, a).
    """

    def __init__(self, n_cls, thr=None, min_areas=(0,), activation='sigmoid',
                 empty_one=True, eps=1e-6, device='cuda', use_lut=True,
                 record_events=False, iou=False):
        self.n_cls = int(n_cls)
        self.thr = ops.threshold_grid(100) if thr is None else \
            np.asarray(thr, dtype=np.float32)
        areas = np.asarray(min_areas, dtype=np.int64)
        if areas.ndim == 1:
            areas = np.broadcast_to(areas, (self.n_cls, areas.shape[0]))
        self.min_areas = np.ascontiguousarray(areas)
        self.activation = activation
        self.empty_one = empty_one
        self.eps = eps
        self.device = torch.device(device)
        # the per-code table only exists for element-wise activations
        self.lut = ops.seg_lut(self.thr, activation, self.device) \
            if use_lut and activation != 'softmax' else None
        b, a = len(self.thr), self.min_areas.shape[1]
        # one buffer for everything that is exchanged / read back: the fp64
        # dice sums (as bit patterns), the int64 counts, the image count
        nd, nm = self.n_cls * b * a, self.n_cls * (1 + 2 * b)
        self.iou = bool(iou)     # also the per-plane (macro) IoU of every cell
        nf = nd * (2 if self.iou else 1)
        self._nd, self._nf = nd, nf
        self.packed = torch.zeros(nf + nm + 1, dtype=torch.int64,
                                  device=self.device)
        self.dice_sum = self.packed[:nd].view(torch.float64) \
            .view(self.n_cls, b, a)
        self.iou_sum = self.packed[nd:nf].view(torch.float64) \
            .view(self.n_cls, b, a) if self.iou else None
        self.micro = self.packed[nf:nf + nm].view(self.n_cls, 1 + 2 * b)
        self.n_img = 0
        self._fresh = False
        self._stage = None
        # (start, end) CUDA events around every histogram launch, for the
        # per-kernel roofline in bench.py
        self.record_events = record_events
        self.kernel_events = []

    def reset(self):
        # the next add() adopts its results instead of accumulating into
        # zeroed tables (no memset / add kernels for a one-chunk pass)
        self._fresh = True
        self.n_img = 0

    # -- device-resident chunk ------------------------------------------------
    def add(self, sources, gt, xforms=None, keep_bins=False):
        """Score one chunk already in HBM; returns its histograms
        [n, C, 2, B+1] (int32) for per-image reporting - with
        ``keep_bins`` the pair (histograms, per-pixel bin indices uint8
        [n, C, H, W]) for ``ops.rle_from_bins`` at the point the search picks
        (``keep_bins=<tensor>``: written into that buffer)."""
        if self.record_events:
            ev = (torch.cuda.Event(enable_timing=True),
                  torch.cuda.Event(enable_timing=True))
            ev[0].record()
        bins = None
        if keep_bins is not False and keep_bins is not None:
            hist, bins = ops.seg_hist(sources, gt, self.thr, xforms=xforms,
                                      activation=self.activation,
                                      bins=keep_bins)
        else:
            hist = ops.seg_hist(sources, gt, self.thr, xforms=xforms,
                                activation=self.activation, lut=self.lut)
        if self.record_events:
            ev[1].record()
            self.kernel_events.append(ev)
        if self.iou:
            dsum, usum, micro = ops.dice_iou_search(
                hist, self.min_areas, self.empty_one, self.eps)
            if self._fresh:
                self.dice_sum.copy_(dsum)
                self.iou_sum.copy_(usum)
                self.micro.copy_(micro)
                self._fresh = False
            else:
                self.dice_sum += dsum
                self.iou_sum += usum
                self.micro += micro
        elif self._fresh:     # first chunk of a pass: written in place
            ops.dice_search(hist, self.min_areas, self.empty_one, self.eps,
                            out=self.dice_sum, micro=self.micro)
            self._fresh = False
        else:
            dsum, micro = ops.dice_search(hist, self.min_areas,
                                          self.empty_one, self.eps)
            self.dice_sum += dsum
            self.micro += micro
        self.n_img += hist.shape[0]
        return (hist, bins) if bins is not None else hist

    # -- host-resident chunks: copies overlapped with compute -----------------
    def _stager(self, arrays, chunk):
        key = tuple((id(a), tuple(a.shape), str(a.dtype)) for a in arrays) + \
            (chunk,)
        if self._stage is None or self._stage[0] != key:
            self._stage = (key, HostStager(arrays, self.device, chunk))
        return self._stage[1]

    def add_host(self, sources, gt, xforms=None, chunk=64, gt_runs=None):
        """Score host tensors (pinned for full speed).  The set is cut into
        chunks of ``chunk`` images; chunk i+1 is copied on a side stream
        while chunk i is scored, through two staging buffers per input
        (``HostStager``; successive calls may follow each other without a
        synchronisation in between).

        ``gt_runs=(runs, off, cnt)`` (host tensors, the layout of
        ``ops.rle_parse`` / ``ops.rle_encode``; ``gt`` is then None): the
        ground truth stays in the run-length form the competition CSVs hold it
        in, is uploaded once (KBs) and decoded per chunk on the device
        (mlc_rle_decode) - 2 instead of 3 bytes per pixel cross PCIe."""
        if isinstance(sources, torch.Tensor):
            sources = [sources]
        sources = list(sources)
        n, c, h, w = sources[0].shape
        runs = off = cnt = None
        if gt_runs is not None:
            if xforms is not None and any(x & ops.MLC_XF_TRANSPOSE
                                          for x in map(_xf_code, xforms)):
                raise ValueError('gt_runs needs the output geometry of the '
                                 'sources: no transposed variants')
            runs, off, cnt = (t.to(self.device, non_blocking=True)
                              for t in gt_runs)
            off, cnt = off.reshape(-1), cnt.reshape(-1)
            arrays = sources
        else:
            arrays = sources + [gt]
        st = self._stager(arrays, chunk)
        parts = [(lo, min(n, lo + chunk)) for lo in range(0, n, chunk)]
        for i, (lo, hi) in enumerate(parts):
            dev = st.fetch(lo, hi)
            if i + 1 < len(parts):
                st.prefetch(*parts[i + 1])
            if gt_runs is not None:
                g = ops.rle_decode(runs, off[lo * c:hi * c],
                                   cnt[lo * c:hi * c], h, w) \
                    .view(hi - lo, c, h, w)
                self.add(dev, g, xforms)
            else:
                self.add(dev[:-1], dev[-1], xforms)
            st.release()

    # -- combine + search -------------------------------------------------------
    def result(self, group=None):
        """Combine ranks (``combine_ranks``) and search the grid."""
        self.packed[-1:].fill_(self.n_img)
        # one collective + one kernel, then one device->host copy for
        # everything the host-side search needs
        packed = combine_packed(self.packed, self._nf, group).cpu().numpy()
        nd, nf = self._nd, self._nf
        n_img = int(packed[-1])
        score = packed[:nd].view(np.float64).reshape(self.dice_sum.shape) / \
            max(n_img, 1)                                     # [C, B, A]
        m = packed[nf:-1].reshape(self.micro.shape)
        b = len(self.thr)
        t_sum, p_sum, i_sum = m[:, 0], m[:, 1:1 + b], m[:, 1 + b:]
        best = []
        for c in range(self.n_cls):
            k, a = divmod(int(np.argmax(score[c])), score.shape[2])
            best.append({'class': c, 'threshold_index': k,
                         'threshold': float(self.thr[k]),
                         'min_area': int(self.min_areas[c, a]),
                         'dice': float(score[c, k, a])})
        union = t_sum[:, None] + p_sum
        extra = {}
        if self.iou:
            # mean over images of the per-plane IoU at every (t, a): not in
            # the reference (builder-defined, see oracle.iou_from_counts)
            extra['iou_score'] = packed[nd:nf].view(np.float64) \
                .reshape(self.dice_sum.shape) / max(n_img, 1)
            extra['iou'] = float(np.mean([
                extra['iou_score'][r['class'], r['threshold_index'],
                                   list(self.min_areas[r['class']])
                                   .index(r['min_area'])] for r in best]))
        return {
            **extra,
            'n_img': int(n_img),
            'score': score,
            'best': best,
            'dice': float(np.mean([r['dice'] for r in best])),
            'micro_T': t_sum, 'micro_P': p_sum, 'micro_I': i_sum,
            'micro_dice': 2.0 * i_sum / (union + self.eps),
            # IoU is not in the reference: self-consistent only
            'micro_iou': i_sum / (union - i_sum + self.eps),
        }


class ClassifyScore:
    """argmax + confusion matrix + accuracy over chunks (config 1):
    worker/reports/classification.py:133-138 and
    examples/digit-recognizer/executors/valid.py:45-52."""

    def __init__(self, n_cls, device='cuda'):
        self.n_cls = int(n_cls)
        self.device = torch.device(device)
        self.cm = torch.zeros((n_cls, n_cls), dtype=torch.int64,
                              device=self.device)
        self.stats = torch.zeros(2, dtype=torch.int64, device=self.device)
        self.n = 0

    def add(self, preds, labels, want_argmax=False):
        _, _, am = ops.confusion(preds, labels, cm=self.cm, stats=self.stats,
                                 want_argmax=want_argmax)
        self.n += preds.shape[0]
        return am

    def result(self, group=None):
        import torch.distributed as dist
        ints = torch.cat([self.cm.reshape(-1), self.stats,
                          torch.tensor([self.n], dtype=torch.int64,
                                       device=self.device)])
        if dist.is_available() and dist.is_initialized() and \
                dist.get_world_size(group) > 1:
            dist.all_reduce(ints, op=dist.ReduceOp.SUM, group=group)
        h = ints.cpu().numpy()
        c = self.n_cls
        cm = h[:c * c].reshape(c, c)
        correct, dropped, n = int(h[c * c]), int(h[c * c + 1]), int(h[-1])
        return {'confusion': cm, 'correct': correct, 'dropped': dropped,
                'n': n, 'accuracy': correct / n if n else float('nan')}
