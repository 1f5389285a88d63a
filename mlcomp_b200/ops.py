"""Torch-tensor front end of the C ABI (include/mlc.h).

Thin by design: validate device / dtype / contiguity, allocate outputs with
torch, pass raw pointers and the current CUDA stream through ctypes.  All
arithmetic happens in the CUDA kernels of ``csrc/``; nothing here computes on
the CPU and nothing falls back to PyTorch ops.
"""
import ctypes
from ctypes import c_float, c_int32, c_void_p

import numpy as np
import torch

from . import _lib
from ._lib import (MLC_ACT_NONE, MLC_ACT_SIGMOID, MLC_F16, MLC_F32, MLC_U8,
                   MLC_XF_HFLIP, MLC_XF_TRANSPOSE, MLC_XF_VFLIP, check)

__all__ = ['MLC_XF_HFLIP', 'MLC_XF_VFLIP', 'MLC_XF_TRANSPOSE',
           'activation_code', 'compose_tta', 'confusion', 'seg_lut',
           'seg_hist', 'seg_prob', 'dice_search', 'dice_iou_search', 'plane_dice', 'rle_encode',
           'threshold_grid', 'counts_from_hist', 'rle_strings',
           'rle_strings_host', 'rle_decode', 'rle_parse', 'mask_composite',
           'rle_from_bins', 'plane_area', 'seg_argmax',
           'mask_outline']


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)


def _stream():
    # (torch.cuda.current_stream() builds a Stream object through several
    # device-index look-ups: 19 us a call, four calls per scoring step)
    if _raw_stream is not None:
        return c_void_p(_raw_stream(torch.cuda.current_device()))
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(t, name, dtypes=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f'{name} must be a CUDA tensor (mlcomp_b200 has no '
                        'CPU path)')
    if not t.is_contiguous():
        raise ValueError(f'{name} must be contiguous')
    if dtypes is not None and t.dtype not in dtypes:
        raise TypeError(f'{name} has dtype {t.dtype}, expected one of {dtypes}')
    return t


def _ptr(t):
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


def _dtype_code(t):
    if t.dtype == torch.float16:
        return MLC_F16
    if t.dtype == torch.float32:
        return MLC_F32
    if t.dtype == torch.uint8:
        return MLC_U8
    raise TypeError(f'unsupported element type {t.dtype}')


def activation_code(activation):
    """utils/torch.py:11-18: falsy -> none, 'sigmoid', 'softmax' (over the
    class dim of every source); anything else raises the reference's
    message."""
    if not activation:
        return MLC_ACT_NONE
    if activation == 'sigmoid':
        return MLC_ACT_SIGMOID
    if activation == 'softmax':
        return _lib.MLC_ACT_SOFTMAX
    raise Exception(f'unknown activation = {activation}')


def compose_tta(tags):
    """Collapse the transform list of one TTA variant to an MLC_XF_* mask
    describing what TtaWrap.inverse (contrib/transform/tta.py:20-31) does:
    'hflip' flips the last dim, 'vflip' the one before, 'transpose' swaps
    them, applied in list order.  Tags are the short names of
    utils/config.py:93-104."""
    tr = vf = hf = False
    for t in tags:
        if t == 'hflip':      # out'[y][x] = out[y][W-1-x]
            if tr:
                vf = not vf
            else:
                hf = not hf
        elif t == 'vflip':
            if tr:
                hf = not hf
            else:
                vf = not vf
        elif t == 'transpose':
            tr = not tr
        else:
            raise Exception(f'Unknwon augmentation {t}')
    return (MLC_XF_HFLIP if hf else 0) | (MLC_XF_VFLIP if vf else 0) | \
        (MLC_XF_TRANSPOSE if tr else 0)


def threshold_grid(bins=100):
    """The search grid: thr[k] = float32(k / bins)."""
    return (np.arange(bins) / bins).astype(np.float32)


# ---------------------------------------------------------------- classify
def confusion(preds, labels, cm=None, stats=None, want_argmax=False):
    """argmax + confusion matrix + hit count (mlc_confusion).

    Returns (cm [C,C] int64, stats [2] int64 = {correct, dropped},
    argmax [n] int32 or None).  Pass cm/stats back in to accumulate chunks.
    """
    lib = _lib.load()
    preds = _need_cuda(preds, 'preds', (torch.float16, torch.float32))
    labels = _need_cuda(labels, 'labels', (torch.int64,))
    if preds.dim() != 2 or labels.dim() != 1 or labels.shape[0] != preds.shape[0]:
        raise ValueError('preds must be [n, C] and labels [n]')
    n, c = preds.shape
    dev = preds.device
    if cm is None:
        cm = torch.zeros((c, c), dtype=torch.int64, device=dev)
    if stats is None:
        stats = torch.zeros(2, dtype=torch.int64, device=dev)
    _need_cuda(cm, 'cm', (torch.int64,))
    _need_cuda(stats, 'stats', (torch.int64,))
    if tuple(cm.shape) != (c, c) or stats.numel() != 2:
        raise ValueError('cm must be [C, C] and stats [2]')
    am = torch.empty(n, dtype=torch.int32, device=dev) if want_argmax else None
    with torch.cuda.device(dev):
        check(lib.mlc_confusion(_ptr(preds), _dtype_code(preds), n, c,
                                _ptr(labels), _ptr(cm), _ptr(stats), _ptr(am),
                                _stream()))
    return cm, stats, am


# ---------------------------------------------------------------- segment
def _sources(sources, xforms, allow_u8=False):
    if isinstance(sources, torch.Tensor):
        sources = [sources]
    sources = list(sources)
    if not sources:
        raise ValueError('at least one source is required')
    if xforms is None:
        xforms = [0] * len(sources)
    xforms = [compose_tta(x) if not isinstance(x, int) else x for x in xforms]
    if len(xforms) != len(sources):
        raise ValueError('one transform per source is required')
    ok = (torch.float16, torch.float32) + ((torch.uint8,) if allow_u8 else ())
    first = _need_cuda(sources[0], 'sources[0]', ok)
    if first.dim() != 4:
        raise ValueError('sources must be [n_img, n_cls, H, W]')
    tr0 = bool(xforms[0] & MLC_XF_TRANSPOSE)
    n, c = first.shape[:2]
    h, w = (first.shape[3], first.shape[2]) if tr0 else first.shape[2:]
    for i, (s, x) in enumerate(zip(sources, xforms)):
        _need_cuda(s, f'sources[{i}]', (first.dtype,))
        want = (n, c, w, h) if x & MLC_XF_TRANSPOSE else (n, c, h, w)
        if tuple(s.shape) != want or s.device != first.device:
            raise ValueError(f'sources[{i}] has shape {tuple(s.shape)}, '
                             f'expected {want} on {first.device}')
    ptrs = (c_void_p * len(sources))(*[s.data_ptr() for s in sources])
    xf = (c_int32 * len(sources))(*xforms)
    return sources, ptrs, xf, int(n), int(c), int(h), int(w)


def _thr(thr):
    thr = np.ascontiguousarray(np.asarray(thr, dtype=np.float32).ravel())
    return thr, thr.ctypes.data_as(ctypes.POINTER(c_float))


def seg_lut(thr, activation='sigmoid', device=None):
    """64 KiB table cnt(activation(x)) for every fp16 code (mlc_seg_lut_build)."""
    lib = _lib.load()
    thr, thr_p = _thr(thr)
    device = torch.device('cuda' if device is None else device)
    lut = torch.empty(_lib.MLC_LUT_BYTES, dtype=torch.uint8, device=device)
    with torch.cuda.device(lut.device):
        check(lib.mlc_seg_lut_build(thr_p, len(thr), activation_code(activation),
                                    _ptr(lut), _stream()))
    return lut


def seg_hist(sources, gt, thr, xforms=None, activation='sigmoid', lut=None,
             out=None, bins=False):
    """Fused TTA/ensemble mean + activation + label-conditioned histogram of
    cnt(p) = #{k: p >= thr[k]} (mlc_seg_hist).  Returns int32
    [n_img, n_cls, 2, B+1] (counts are < 2**31).

    ``bins=True`` (mlc_seg_hist_bins) also returns the bin index of every
    pixel, uint8 [n_img, n_cls, H, W]: ``bins > k`` is the mask at thr[k]
    (``rle_from_bins``), 1 byte per pixel instead of the sources again.
    ``bins=<uint8 tensor of that shape>`` writes them there (a loop over
    chunks then allocates nothing: at 10 GB per buffer a cudaMalloc or
    cudaFree of the caching allocator stalls the step by tens of ms)."""
    lib = _lib.load()
    sources, ptrs, xf, n, c, h, w = _sources(sources, xforms)
    gt = _need_cuda(gt, 'gt', (torch.uint8,))
    if tuple(gt.shape) != (n, c, h, w):
        raise ValueError(f'gt has shape {tuple(gt.shape)}, expected '
                         f'{(n, c, h, w)} uint8 one-hot planes')
    thr, thr_p = _thr(thr)
    dev = sources[0].device
    if out is None:
        out = torch.empty((n, c, 2, len(thr) + 1), dtype=torch.int32, device=dev)
    _need_cuda(out, 'out', (torch.int32,))
    if out.numel() != n * c * 2 * (len(thr) + 1):
        raise ValueError('out has the wrong size')
    if lut is not None:
        _need_cuda(lut, 'lut', (torch.uint8,))
        if lut.numel() != _lib.MLC_LUT_BYTES:
            raise ValueError('lut must come from seg_lut()')
    if bins is not False and bins is not None:
        if len(thr) > 255:
            raise ValueError('bins need n_thr <= 255')
        if isinstance(bins, torch.Tensor):
            bins_out = _need_cuda(bins, 'bins', (torch.uint8,))
            if tuple(bins_out.shape) != (n, c, h, w) or \
                    not bins_out.is_contiguous():
                raise ValueError(f'bins has shape {tuple(bins_out.shape)}, '
                                 f'expected contiguous {(n, c, h, w)} uint8')
        else:
            bins_out = torch.empty((n, c, h, w), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            check(lib.mlc_seg_hist_bins(
                ptrs, xf, len(sources), _dtype_code(sources[0]), _ptr(gt), n,
                c, h, w, thr_p, len(thr), activation_code(activation),
                _ptr(out), _ptr(bins_out), _stream()))
        return out, bins_out
    with torch.cuda.device(dev):
        check(lib.mlc_seg_hist(ptrs, xf, len(sources), _dtype_code(sources[0]),
                               _ptr(gt), n, c, h, w, thr_p, len(thr),
                               activation_code(activation), _ptr(lut),
                               _ptr(out), _stream()))
    return out


def rle_from_bins(bins, k_per_class, min_area_per_class=0, capacity=None,
                  hist=None):
    """Run-length pairs of the masks (bins > k_c) = (prob >= thr[k_c]) from
    the per-pixel bin indices ``seg_hist(..., bins=True)`` kept (same return
    values as ``rle_encode``).  ``hist``: the histograms of the same pass -
    planes whose mask is empty or below the min-area are then not read."""
    k = np.asarray(k_per_class, dtype=np.float32) + np.float32(0.5)
    known = plane_area(hist, k_per_class) if hist is not None else None
    return rle_encode([bins], k, min_area_per_class, activation=None,
                      capacity=capacity, known_area=known)


def seg_prob(sources, xforms=None, activation='sigmoid'):
    """Fused mean of activated, un-transformed sources -> fp32 [n,c,H,W]."""
    lib = _lib.load()
    sources, ptrs, xf, n, c, h, w = _sources(sources, xforms)
    dev = sources[0].device
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.mlc_seg_prob(ptrs, xf, len(sources), _dtype_code(sources[0]),
                               n, c, h, w, activation_code(activation),
                               _ptr(out), _stream()))
    return out


def seg_argmax(sources, xforms=None, activation='softmax', gt_onehot=None,
               gt_label=None, want_label=True, want_onehot=False,
               want_counts=False, want_confusion=False):
    """Per-pixel arg-max over the classes of the fused scores
    (mlc_seg_argmax): first maximal class, NaN maximal, ties -> lowest.
    Returns a dict with the requested outputs: 'label' uint8 [n, H, W],
    'onehot' uint8 [n, C, H, W], 'counts' int64 [n, C, 3] = {T, P, I} of the
    one-hot masks, 'confusion' int64 [C, C] (rows = truth; needs gt_label)."""
    lib = _lib.load()
    sources, ptrs, xf, n, c, h, w = _sources(sources, xforms)
    dev = sources[0].device
    if c > 255:
        raise ValueError('at most 255 classes')
    if gt_onehot is not None:
        _need_cuda(gt_onehot, 'gt_onehot', (torch.uint8,))
        if tuple(gt_onehot.shape) != (n, c, h, w):
            raise ValueError(f'gt_onehot must be {(n, c, h, w)} uint8')
    if gt_label is not None:
        _need_cuda(gt_label, 'gt_label', (torch.uint8,))
        if tuple(gt_label.shape) != (n, h, w):
            raise ValueError(f'gt_label must be {(n, h, w)} uint8')
    out = {}
    if want_label:
        out['label'] = torch.empty((n, h, w), dtype=torch.uint8, device=dev)
    if want_onehot:
        out['onehot'] = torch.empty((n, c, h, w), dtype=torch.uint8,
                                    device=dev)
    if want_counts:
        out['counts'] = torch.empty((n, c, 3), dtype=torch.int64, device=dev)
    if want_confusion:
        out['confusion'] = torch.empty((c, c), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        check(lib.mlc_seg_argmax(
            ptrs, xf, len(sources), _dtype_code(sources[0]), n, c, h, w,
            activation_code(activation), _ptr(gt_onehot), _ptr(gt_label),
            _ptr(out.get('label')), _ptr(out.get('onehot')),
            _ptr(out.get('counts')), _ptr(out.get('confusion')), _stream()))
    return out


def _areas(min_areas, n_cls):
    a = np.asarray(min_areas, dtype=np.int64)
    if a.ndim == 1:
        a = np.broadcast_to(a, (n_cls, a.shape[0]))
    if a.ndim != 2 or a.shape[0] != n_cls:
        raise ValueError('min_areas must be [A] or [n_cls, A]')
    a = np.ascontiguousarray(a.astype(np.int32))
    return a, a.ctypes.data_as(ctypes.POINTER(c_int32))


def dice_iou_search(hist, min_areas=(0,), empty_one=True, eps=1e-6):
    """``dice_search`` plus the IoU sums of the same masks
    (mlc_dice_iou_search): (dice_sum, iou_sum fp64 [C,B,A], micro)."""
    lib = _lib.load()
    hist = _need_cuda(hist, 'hist', (torch.int32,))
    n, c, _, nb = hist.shape
    b = nb - 1
    areas, areas_p = _areas(min_areas, c)
    a = areas.shape[1]
    dev = hist.device
    dice = torch.empty((c, b, a), dtype=torch.float64, device=dev)
    iou = torch.empty((c, b, a), dtype=torch.float64, device=dev)
    micro = torch.empty((c, 1 + 2 * b), dtype=torch.int64, device=dev)
    need = 2 * lib.mlc_dice_search_workspace(n, c, b, a)
    work = torch.empty(max(need, 8), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.mlc_dice_iou_search(
            _ptr(hist), n, c, b, areas_p, a, int(bool(empty_one)), float(eps),
            _ptr(dice), _ptr(iou), _ptr(micro), _ptr(work), work.numel(),
            _stream()))
    return dice, iou, micro


def dice_search(hist, min_areas=(0,), empty_one=True, eps=1e-6,
                want_micro=True, out=None, micro=None):
    """Histograms -> sum over images of dice for every (class, threshold,
    min-area) (mlc_dice_search).  Returns (dice_sum fp64 [C,B,A],
    micro int64 [C, 1+2B] = {T, P[k], I[k]} or None); ``out`` / ``micro``:
    contiguous device tensors of those shapes to write into."""
    lib = _lib.load()
    hist = _need_cuda(hist, 'hist', (torch.int32,))
    if hist.dim() != 4 or hist.shape[2] != 2:
        raise ValueError('hist must be [n_img, n_cls, 2, B+1]')
    n, c, _, nb = hist.shape
    b = nb - 1
    areas, areas_p = _areas(min_areas, c)
    a = areas.shape[1]
    dev = hist.device
    if out is None:
        out = torch.empty((c, b, a), dtype=torch.float64, device=dev)
    if micro is None and want_micro:
        micro = torch.empty((c, 1 + 2 * b), dtype=torch.int64, device=dev)
    _need_cuda(out, 'out', (torch.float64,))
    if out.numel() != c * b * a or \
            (micro is not None and micro.numel() != c * (1 + 2 * b)):
        raise ValueError('out / micro have the wrong size')
    need = lib.mlc_dice_search_workspace(n, c, b, a)
    work = torch.empty(max(need, 8), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.mlc_dice_search(_ptr(hist), n, c, b, areas_p, a,
                                  int(bool(empty_one)), float(eps), _ptr(out),
                                  _ptr(micro), _ptr(work), work.numel(),
                                  _stream()))
    return out, micro


def plane_dice(hist, k_per_class, min_area_per_class, empty_one=True,
               eps=1e-6):
    """Per-plane dice at one (threshold index, min_area) per class
    (mlc_plane_dice): fp64 [n_img, n_cls]."""
    lib = _lib.load()
    hist = _need_cuda(hist, 'hist', (torch.int32,))
    n, c, _, nb = hist.shape
    k = np.ascontiguousarray(np.broadcast_to(
        np.asarray(k_per_class, dtype=np.int32), (c,)))
    ar = np.ascontiguousarray(np.broadcast_to(
        np.asarray(min_area_per_class, dtype=np.int32), (c,)))
    out = torch.empty((n, c), dtype=torch.float64, device=hist.device)
    with torch.cuda.device(hist.device):
        check(lib.mlc_plane_dice(_ptr(hist), n, c, nb - 1,
                                 k.ctypes.data_as(ctypes.POINTER(c_int32)),
                                 ar.ctypes.data_as(ctypes.POINTER(c_int32)),
                                 int(bool(empty_one)), float(eps), _ptr(out),
                                 _stream()))
    return out


def plane_area(hist, k_per_class):
    """Mask area of every plane at thr[k_c] (mlc_plane_area): int32 [n, C]."""
    lib = _lib.load()
    hist = _need_cuda(hist, 'hist', (torch.int32,))
    n, c, _, nb = hist.shape
    k = np.ascontiguousarray(np.broadcast_to(
        np.asarray(k_per_class, dtype=np.int32), (c,)))
    out = torch.empty((n, c), dtype=torch.int32, device=hist.device)
    with torch.cuda.device(hist.device):
        check(lib.mlc_plane_area(_ptr(hist), n, c, nb - 1,
                                 k.ctypes.data_as(ctypes.POINTER(c_int32)),
                                 _ptr(out), _stream()))
    return out


def counts_from_hist(hist):
    """Host-side view helper for tests/reports: (T [n,c], P [n,c,B],
    I [n,c,B]) int64 from a histogram tensor, via torch integer ops on the
    device (not on the timed path)."""
    h = hist.to(torch.int64)
    tot = h[:, :, 0] + h[:, :, 1]
    p = tot.flip(-1).cumsum(-1).flip(-1)[..., 1:]
    i = h[:, :, 1].flip(-1).cumsum(-1).flip(-1)[..., 1:]
    return h[:, :, 1].sum(-1), p, i


# -------------------------------------------------------------------- rle
def rle_encode(sources, thr_per_class, min_area_per_class=0, xforms=None,
               activation='sigmoid', capacity=None, known_area=None):
    """(prob >= thr_c) + min-area reject + column-major run-length pairs
    (mlc_rle_encode).  Returns device tensors
    (runs int32 [total], off int64 [n,c], cnt int32 [n,c], area int32 [n,c]);
    plane p owns runs[off[p] : off[p] + 2*cnt[p]] as (start, length) pairs.
    Synchronises once to learn the total size (and retries if the first
    guess of the capacity was too small).

    ``known_area`` (int32 [n, c] on the device, e.g. ``plane_area(hist, k)``
    when thr_c = thr_grid[k_c]): the exact mask areas; planes that end up
    without runs are then not read at all (mlc_rle_encode_known)."""
    lib = _lib.load()
    sources, ptrs, xf, n, c, h, w = _sources(sources, xforms, allow_u8=True)
    dev = sources[0].device
    thr = np.ascontiguousarray(np.broadcast_to(
        np.asarray(thr_per_class, dtype=np.float32), (c,)))
    ar = np.ascontiguousarray(np.broadcast_to(
        np.asarray(min_area_per_class, dtype=np.int32), (c,)))
    off = torch.empty((n, c), dtype=torch.int64, device=dev)
    cnt = torch.empty((n, c), dtype=torch.int32, device=dev)
    area = torch.empty((n, c), dtype=torch.int32, device=dev)
    total = torch.empty(1, dtype=torch.int64, device=dev)  # cleared by the call
    if known_area is not None:
        known_area = _need_cuda(known_area, 'known_area', (torch.int32,))
        if known_area.numel() != n * c:
            raise ValueError('known_area must be [n_img, n_cls]')
    if capacity is None:
        # 1 run pair per 64 px is generous for real masks; retried if not
        capacity = max(1024, n * c * h * w // 32)
    while True:
        runs = torch.empty(capacity, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            if known_area is None:
                check(lib.mlc_rle_encode(
                    ptrs, xf, len(sources), _dtype_code(sources[0]), n, c, h,
                    w, thr.ctypes.data_as(ctypes.POINTER(c_float)),
                    ar.ctypes.data_as(ctypes.POINTER(c_int32)),
                    activation_code(activation), _ptr(runs), capacity,
                    _ptr(off), _ptr(cnt), _ptr(area), _ptr(total), _stream()))
            else:
                check(lib.mlc_rle_encode_known(
                    ptrs, xf, len(sources), _dtype_code(sources[0]), n, c, h,
                    w, thr.ctypes.data_as(ctypes.POINTER(c_float)),
                    ar.ctypes.data_as(ctypes.POINTER(c_int32)),
                    activation_code(activation), _ptr(known_area),
                    _ptr(runs), capacity, _ptr(off), _ptr(cnt), _ptr(area),
                    _ptr(total), _stream()))
        need = int(total.item())
        if need <= capacity:
            return runs[:need], off, cnt, area
        capacity = need


def rle_strings(runs, off, cnt):
    """The reference's strings (contrib/transform/rle.py:13:
    ' '.join(str(x) for x in runs)), one per plane, '' for an empty mask.
    The decimal text is produced on the device (mlc_rle_format); the host
    only slices one byte buffer.  Synchronises (it returns Python strings)."""
    lib = _lib.load()
    dev = off.device
    planes = off.numel()
    str_off = torch.empty(planes, dtype=torch.int64, device=dev)
    str_len = torch.empty(planes, dtype=torch.int32, device=dev)
    total = torch.empty(1, dtype=torch.int64, device=dev)  # cleared by the call
    # a value is at most 10 digits + 1 blank: 11 characters per int32 always
    # suffice, so the size never has to be asked for before allocating
    capacity = max(64, 11 * int(runs.numel()))
    chars = torch.empty(capacity, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.mlc_rle_format(_ptr(runs), _ptr(off), _ptr(cnt), planes,
                                 _ptr(chars), capacity, _ptr(str_off),
                                 _ptr(str_len), _ptr(total), _stream()))
    # two device->host copies in all: the index (offsets, lengths, total) in
    # one packed tensor, then exactly the characters that were written
    idx = torch.cat([str_off, str_len.to(torch.int64), total]).cpu().numpy()
    o, n, need = idx[:planes], idx[planes:2 * planes], int(idx[-1])
    text = chars[:need].cpu().numpy().tobytes()
    return [text[a:a + b].decode('ascii') for a, b in zip(o, n)]


def rle_strings_host(runs, off, cnt):
    """Host-side formatting of the same integers (used by the tests to check
    the device formatter)."""
    runs_h = runs.cpu().numpy()
    off_h = off.cpu().numpy().ravel()
    cnt_h = cnt.cpu().numpy().ravel()
    return [' '.join(map(str, runs_h[o:o + 2 * k].tolist())) if k else ''
            for o, k in zip(off_h, cnt_h)]


def rle_parse(strings, device='cuda'):
    """Host side of decoding: split the reference's "start length ..."
    strings (rle.py:22-26) into one int32 pair buffer + per-plane offsets /
    counts on the device."""
    arrays = [np.array(t.split(), dtype=np.int32) if t else
              np.zeros(0, np.int32) for t in strings]
    for a in arrays:
        if a.size % 2:
            raise ValueError('an RLE string needs an even number of values')
        if a.size and (a[0::2].min() < 1 or a[1::2].min() < 0):
            raise ValueError('RLE starts are 1-based and lengths are not '
                             'negative (rle.py:22-26)')
    cnt = np.array([a.size // 2 for a in arrays], dtype=np.int32)
    off = np.concatenate([[0], np.cumsum(2 * cnt.astype(np.int64))[:-1]]) \
        if len(arrays) else np.zeros(0, np.int64)
    flat = np.concatenate(arrays) if arrays else np.zeros(0, np.int32)
    dev = torch.device(device)
    return (torch.from_numpy(np.ascontiguousarray(flat)).to(dev),
            torch.from_numpy(np.ascontiguousarray(off, dtype=np.int64)).to(dev),
            torch.from_numpy(cnt).to(dev))


def rle_decode(runs, off, cnt, h, w):
    """(start, length) pairs -> uint8 masks [planes, h, w] on the device
    (mlc_rle_decode): rle2mask for a whole set in one launch."""
    lib = _lib.load()
    dev = off.device
    planes = off.numel()
    if runs.numel() == 0:
        runs = torch.zeros(2, dtype=torch.int32, device=dev)
    out = torch.empty((planes, h, w), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.mlc_rle_decode(_ptr(runs.contiguous()), _ptr(off.contiguous()),
                                 _ptr(cnt.contiguous()), planes, h, w,
                                 _ptr(out), _stream()))
    return out


def _palette(colors, n_cls, wrap):
    """[n_cls][3] uint8 host array for the C ABI, or None for the reference's
    default colour.  ``wrap``: numpy's modulo-256 of encode_pred's uint8
    product; otherwise OpenCV's saturation of a colour scalar."""
    if colors is None:
        return None, None
    c = np.asarray(colors, dtype=np.int64).reshape(-1, 3)[:n_cls]
    if c.shape[0] < n_cls:
        raise IndexError('list index out of range')   # colors[i], as there
    c = (c % 256) if wrap else np.clip(c, 0, 255)
    c = np.ascontiguousarray(c, dtype=np.uint8)
    return c, c.ctypes.data_as(c_void_p)


def mask_composite(mask, colors=None):
    """encode_pred (worker/reports/segmenation.py:71-80) for a batch:
    mask [n, C, H, W] uint8 on the device -> [n, H, W, 3] uint8 colour
    composite (mlc_mask_composite).  ``colors`` None = white."""
    lib = _lib.load()
    _need_cuda(mask, 'mask', (torch.uint8,))
    if mask.dim() != 4:
        raise ValueError('mask must be [n, C, H, W]')
    mask = mask.contiguous()
    n, c, h, w = mask.shape
    keep, ptr = _palette(colors, c, wrap=True)
    out = torch.empty((n, h, w, 3), dtype=torch.uint8, device=mask.device)
    with torch.cuda.device(mask.device):
        check(lib.mlc_mask_composite(_ptr(mask), n, c, h, w, ptr, _ptr(out),
                                     _stream()))
    return out


def mask_outline(img, mask, colors=None):
    """plot_mask (worker/reports/segmenation.py:82-96) for a batch: draws the
    outline of every class plane of mask [n, C, H, W] uint8 over
    img [n, H, W, 3] uint8 IN PLACE (mlc_mask_outline) and returns img.
    ``colors`` None (or empty, as the reference tests truthiness) = green."""
    lib = _lib.load()
    _need_cuda(img, 'img', (torch.uint8,))
    _need_cuda(mask, 'mask', (torch.uint8,))
    if mask.dim() != 4 or img.dim() != 4 or img.shape[3] != 3 or \
            tuple(img.shape[:3]) != (mask.shape[0],) + tuple(mask.shape[2:]):
        raise ValueError('img must be [n, H, W, 3] and mask [n, C, H, W]')
    if not img.is_contiguous():
        raise ValueError('img must be contiguous (it is updated in place)')
    mask = mask.contiguous()
    n, c, h, w = mask.shape
    keep, ptr = _palette(colors if colors else None, c, wrap=False)
    with torch.cuda.device(mask.device):
        check(lib.mlc_mask_outline(_ptr(img), _ptr(mask), n, c, h, w, ptr,
                                   _stream()))
    return img
