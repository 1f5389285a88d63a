"""NumPy / CPU-torch restatement of the reference's scoring hot path.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``): never imported by the
product package.  Every function names the reference file:line it follows
(paths relative to ``/root/reference``).  Where the reference has no code
(threshold grid, TTA/ensemble mean, dataset aggregation: SURVEY.md section 0
fact 4) the function says "builder-defined" and fixes the arithmetic that the
CUDA path must reproduce.
"""
from __future__ import annotations

import numpy as np
import torch

__all__ = [
    'plane_dice', 'rle_encode', 'rle_decode', 'flip_axis', 'tta_inverse',
    'activate', 'confusion', 'confusion_numpy',
    'classify_scores', 'classify_score_final', 'threshold_grid',
    'fuse_probs', 'plane_counts', 'bin_counts', 'dice_from_counts',
    'dice_grid_slow', 'dice_grid_fast', 'grid_scores', 'best_point',
    'micro_stats', 'rle_submission', 'encode_pred', 'plot_mask', 'TTA_TAGS',
    'seg_argmax', 'onehot_counts', 'iou_from_counts',
]

TTA_TAGS = ('hflip', 'vflip', 'transpose')


# --------------------------------------------------------------------------
# A6  dice_numpy                          mlcomp/contrib/metrics/dice.py:4-23
# --------------------------------------------------------------------------
def plane_dice(gt, prob, threshold=None, min_area=None, empty_one=True,
               eps=1e-6):
    """Dice of ONE (H, W) plane at ONE (threshold, min_area) point.

    Follows dice.py:4-23 step by step, with the same NumPy calls so that the
    dtypes (fp32 target sum, uint64 prediction sum, fp64 quotient) and the
    ``int 1`` early return are the reference's own.
    """
    pred = prob
    if threshold is not None:                       # dice.py:6-8
        pred = (prob >= threshold).astype(np.uint8)
    n_gt = gt.sum()                                 # dice.py:10
    n_pred = pred.sum()                             # dice.py:11
    if min_area and n_pred < min_area:              # dice.py:13-15
        pred = np.zeros(pred.shape, dtype=np.uint8)
        n_pred = 0
    if empty_one and n_gt == 0 and n_pred == 0:     # dice.py:17-18
        return 1
    inter = (gt * pred).sum()                       # dice.py:20
    return 2 * inter / (n_gt + n_pred + eps)        # dice.py:21-23


# --------------------------------------------------------------------------
# A7 / A8  mask2rle, rle2mask             mlcomp/contrib/transform/rle.py:4-31
# --------------------------------------------------------------------------
def rle_encode(mask) -> str:
    """Kaggle run-length string of a (H, W) {0,1} mask (rle.py:4-13).

    Column-major scan (``mask.T`` flattened), 1-based starts, "start length"
    pairs joined by single blanks; '' for an empty mask.
    """
    flat = np.asarray(mask).T.flatten()                       # rle.py:9
    padded = np.concatenate([[0], flat, [0]])                 # rle.py:10
    edges = np.where(padded[1:] != padded[:-1])[0] + 1        # rle.py:11
    edges[1::2] -= edges[::2]                                 # rle.py:12
    return ' '.join(str(v) for v in edges)                    # rle.py:13


def rle_decode(text: str, shape):
    """Inverse of :func:`rle_encode`; ``shape`` is (width, height) and the
    result is (height, width) uint8 (rle.py:16-31)."""
    tok = text.split()
    starts = np.asarray(tok[0::2], dtype=int) - 1             # rle.py:24-26
    lengths = np.asarray(tok[1::2], dtype=int)
    flat = np.zeros(shape[0] * shape[1], dtype=np.uint8)      # rle.py:28
    for s, n in zip(starts, lengths):                         # rle.py:29-30
        flat[s:s + n] = 1
    return flat.reshape(shape).T                              # rle.py:31


# --------------------------------------------------------------------------
# A4  flip                                mlcomp/contrib/torch/tensors.py:4-8
# --------------------------------------------------------------------------
def flip_axis(x: torch.Tensor, dim: int) -> torch.Tensor:
    """Reverse ``x`` along ``dim`` by advanced indexing with a descending
    ``arange`` (tensors.py:4-8)."""
    sel = [slice(None)] * x.dim()
    sel[dim] = torch.arange(x.size(dim) - 1, -1, -1, dtype=torch.long,
                            device=x.device)
    return x[tuple(sel)]


# --------------------------------------------------------------------------
# A3 / A5  TtaWrap.inverse, parse_albu_short
#          mlcomp/contrib/transform/tta.py:20-31, mlcomp/utils/config.py:93-104
# --------------------------------------------------------------------------
def tta_inverse(a: torch.Tensor, tags) -> torch.Tensor:
    """Undo the test-time transforms of one variant, in list order.

    tta.py:20-31 dispatches on albumentations classes; albumentations is not
    in this image, so the classes are replaced by the short tags that
    config.py:95-100 maps onto them: 'hflip' -> HorizontalFlip (flip last
    dim), 'vflip' -> VerticalFlip (flip last-1), 'transpose' -> Transpose
    (permute the two spatial dims).  Any other tag raises, as
    config.py:102 does.
    """
    last = a.dim() - 1                                        # tta.py:21
    for t in tags:                                            # tta.py:22
        if t == 'hflip':
            a = flip_axis(a, last)                            # tta.py:23-24
        elif t == 'vflip':
            a = flip_axis(a, last - 1)                        # tta.py:25-26
        elif t == 'transpose':
            perm = (0, 1, 3, 2) if a.dim() == 4 else (0, 2, 1)  # tta.py:28
            a = a.permute(*perm)                              # tta.py:29
        else:
            raise Exception(f'Unknwon augmentation {t}')      # config.py:102
    return a


# --------------------------------------------------------------------------
# A1  apply_activation                    mlcomp/utils/torch.py:11-18
# --------------------------------------------------------------------------
def activate(x: torch.Tensor, activation):
    """None/'' -> x; 'sigmoid' -> torch.sigmoid; 'softmax' -> softmax over
    dim 1; anything else raises (utils/torch.py:11-18).  fp16 input is
    up-cast to fp32 first: the reference's models emit fp32 and SURVEY.md
    section 7 fixes "fp16 logits are up-cast before the activation"."""
    if x.dtype == torch.float16:
        x = x.float()
    if not activation:
        return x
    if activation == 'sigmoid':
        return torch.sigmoid(x)
    if activation == 'softmax':
        return torch.softmax(x, 1)
    raise Exception(f'unknown activation = {activation}')


# --------------------------------------------------------------------------
# A9  confusion matrix      mlcomp/worker/reports/classification.py:133-138
# --------------------------------------------------------------------------
def confusion(targets, preds):
    """The exact sklearn call of classification.py:133-138: rows = truth,
    columns = ``preds.argmax(axis=1)``, labels = arange(C); int64."""
    from sklearn.metrics import confusion_matrix
    preds = np.asarray(preds)
    m = confusion_matrix(targets, preds.argmax(axis=1),
                         labels=np.arange(preds.shape[1]))
    return np.array(m)


def confusion_numpy(targets, preds):
    """sklearn-free restatement of :func:`confusion` (scikit-learn is an
    unpinned third-party dependency, requirements.txt:18): pairs whose truth
    lies outside [0, C) are dropped, as sklearn does for labels not listed."""
    preds = np.asarray(preds)
    c = preds.shape[1]
    y = np.asarray(targets).astype(np.int64)
    p = preds.argmax(axis=1).astype(np.int64)
    ok = (y >= 0) & (y < c)
    return np.bincount(y[ok] * c + p[ok], minlength=c * c).reshape(c, c) \
        .astype(np.int64)


# --------------------------------------------------------------------------
# A12  ValidMnist.score / score_final
#      examples/digit-recognizer/executors/valid.py:45-52
# --------------------------------------------------------------------------
def classify_scores(preds, y):
    """valid.py:47 with ``float`` for the removed ``np.float``."""
    return (np.asarray(preds).argmax(axis=1) == np.asarray(y)).astype(float)


def classify_score_final(all_scores):
    """valid.py:52."""
    return np.mean(all_scores)


# --------------------------------------------------------------------------
# Builder-defined pieces (no reference code: SURVEY.md section 8 D)
# --------------------------------------------------------------------------
def threshold_grid(bins: int = 100) -> np.ndarray:
    """fp32 table thr[k] = float32(k / bins); comparisons are fp32 >=."""
    return (np.arange(bins) / bins).astype(np.float32)


def fuse_probs(sources, tags_per_source, activation='sigmoid') -> np.ndarray:
    """Per-source: activation, then TtaWrap.inverse, then .numpy()
    (the order of utils/torch.py:24-29), then the builder-defined mean over
    sources: fp32, summed left to right, divided once by fp32(S)."""
    acc = None
    for src, tags in zip(sources, tags_per_source):
        p = activate(torch.as_tensor(src), activation)        # torch.py:25
        p = tta_inverse(p, tags)                              # torch.py:27-28
        p = p.detach().cpu().numpy().astype(np.float32)       # torch.py:29
        acc = p.copy() if acc is None else acc + p
    return acc / np.float32(len(sources))


def seg_argmax(sources, tags_per_source, activation='softmax') -> np.ndarray:
    """Arg-max masks of the multiclass path: apply_activation(x, 'softmax')
    per source (utils/torch.py:16-17), TtaWrap.inverse, the mean over sources
    (:func:`fuse_probs`), then numpy's arg-max over the class axis, the call
    the reference makes at worker/reports/classification.py:111,136 (first
    maximal index; a NaN counts as maximal).  int64 [N, H, W]."""
    return np.argmax(fuse_probs(sources, tags_per_source, activation), axis=1)


def onehot_counts(label, gt_onehot):
    """{T, P, I} of dice.py:10-11,20 for the one-hot masks (label == c):
    int64 [N, C, 3]."""
    n, c = gt_onehot.shape[:2]
    out = np.zeros((n, c, 3), np.int64)
    for i in range(n):
        for j in range(c):
            m = label[i] == j
            g = gt_onehot[i, j] != 0
            out[i, j] = (g.sum(), m.sum(), (m & g).sum())
    return out


def bin_counts(gt, prob, thr):
    """Label-conditioned histogram of cnt(p) = #{k: p >= thr[k]} for one
    plane: int64 [2, B+1].  thr must be ascending.  NaN lands in bin 0."""
    thr = np.asarray(thr, dtype=np.float32)
    prob = np.asarray(prob, dtype=np.float32)
    cnt = np.searchsorted(thr, prob.ravel(), side='right')
    cnt[np.isnan(prob.ravel())] = 0
    lab = (np.asarray(gt).ravel() != 0).astype(np.int64)
    nb = thr.size + 1
    return np.bincount(lab * nb + cnt, minlength=2 * nb).reshape(2, nb)


def plane_counts(gt, prob, thr):
    """Integers behind dice.py:10-11,20 for every threshold of one plane:
    T = gt.sum(), P[k] = (prob >= thr[k]).sum(), I[k] = (gt*pred).sum()."""
    h = bin_counts(gt, prob, thr)
    tot = h[0] + h[1]
    # P[k] = number of pixels with cnt > k
    p = tot[::-1].cumsum()[::-1][1:]
    i = h[1][::-1].cumsum()[::-1][1:]
    return int(h[1].sum()), p.astype(np.int64), i.astype(np.int64)


def dice_from_counts(t, p, i, min_area=None, empty_one=True, eps=1e-6):
    """dice.py:13-23 evaluated from the exact integers (scalar or arrays).
    Bit-equal to :func:`plane_dice` because fp32 sums of < 2**24 ones are
    exact and the final divide is a single fp64 operation."""
    p = np.asarray(p, dtype=np.int64)
    i = np.asarray(i, dtype=np.int64)
    if min_area:
        small = p < min_area
        p = np.where(small, 0, p)
        i = np.where(small, 0, i)
    d = 2.0 * i / (np.float64(t) + p + eps)
    if empty_one and t == 0:
        d = np.where(p == 0, 1.0, d)
    return d


def iou_from_counts(t, p, i, min_area=None, empty_one=True, eps=1e-6):
    """BUILDER-DEFINED, UNPINNED: the reference has no IoU
    (contrib/metrics/dice.py:4-23 is dice only).  Same masks
    and the same two rules as :func:`dice_from_counts` - the min-area rule
    empties the prediction (dice.py:13-15), empty truth and prediction score 1
    (dice.py:17-18) - with I / (T + P - I + eps) in place of 2I / (T + P +
    eps).  Self-consistent only: it checks the kernel against this formula,
    not against the reference."""
    p = np.asarray(p, dtype=np.int64)
    i = np.asarray(i, dtype=np.int64)
    if min_area:
        small = p < min_area
        p = np.where(small, 0, p)
        i = np.where(small, 0, i)
    d = i / (np.float64(t) + p - i + eps)
    if empty_one and t == 0:
        d = np.where(p == 0, 1.0, d)
    return d


def dice_grid_slow(gt, prob, thr, min_areas, empty_one=True, eps=1e-6):
    """[N, C, B, A] fp64 by calling :func:`plane_dice` at every grid point -
    the loop the missing ``Valid.score`` ran.  ``gt`` is the
    reference's fp32 one-hot (contrib/dataset/segment.py:36-56)."""
    n, c = prob.shape[:2]
    out = np.zeros((n, c, len(thr), len(min_areas)))
    for a in range(n):
        for b in range(c):
            g = gt[a, b].astype(np.float32)
            for k, t in enumerate(thr):
                for m, area in enumerate(min_areas):
                    out[a, b, k, m] = plane_dice(
                        g, prob[a, b], threshold=np.float32(t),
                        min_area=int(area), empty_one=empty_one, eps=eps)
    return out


def dice_grid_fast(gt, prob, thr, min_areas, empty_one=True, eps=1e-6):
    """Same table as :func:`dice_grid_slow` through the integer counts."""
    n, c = prob.shape[:2]
    out = np.zeros((n, c, len(thr), len(min_areas)))
    for a in range(n):
        for b in range(c):
            t, p, i = plane_counts(gt[a, b], prob[a, b], thr)
            for m, area in enumerate(min_areas):
                out[a, b, :, m] = dice_from_counts(t, p, i, int(area),
                                                   empty_one, eps)
    return out


def grid_scores(dice_table):
    """score[c, t, a] = mean over images (builder-defined)."""
    return dice_table.mean(axis=0)


def best_point(scores):
    """First arg-max over (t, a) per class: [(k, m), ...]."""
    res = []
    for c in range(scores.shape[0]):
        flat = int(np.argmax(scores[c]))
        res.append(divmod(flat, scores.shape[2]))
    return res


def micro_stats(gt, prob, thr):
    """Global per-class integers: T[c], P[c, k], I[c, k] summed over images
    (what the NCCL all-reduce carries), and micro dice / IoU from them."""
    n, c = prob.shape[:2]
    b = len(thr)
    tt = np.zeros(c, np.int64)
    pp = np.zeros((c, b), np.int64)
    ii = np.zeros((c, b), np.int64)
    for a in range(n):
        for j in range(c):
            t, p, i = plane_counts(gt[a, j], prob[a, j], thr)
            tt[j] += t
            pp[j] += p
            ii[j] += i
    return tt, pp, ii


def rle_submission(prob_plane, threshold, min_area=0) -> str:
    """Builder-defined Infer.submit leaf: binarise with
    fp32 >=, zero the mask when its area is below ``min_area`` (the rule of
    dice.py:13-15), then rle.py:4-13."""
    m = (np.asarray(prob_plane, np.float32) >= np.float32(threshold)) \
        .astype(np.uint8)
    if min_area and m.sum() < min_area:
        m[:] = 0
    return rle_encode(m)


# --------------------------------------------------------------------------
# A11  encode_pred              mlcomp/worker/reports/segmenation.py:71-80
# --------------------------------------------------------------------------
def encode_pred(mask, colors=None):
    """(C, H, W) {0,1} -> (H, W, 3) uint8 colour composite; white when no
    palette (segmenation.py:71-80).  uint8 addition wraps, as there."""
    res = np.zeros((*mask.shape[1:], 3), dtype=np.uint8)
    for idx, plane in enumerate(mask):
        rgb = np.repeat(plane[:, :, None], 3, axis=2)
        col = colors[idx] if colors is not None else (255, 255, 255)
        res += (rgb * col).astype(np.uint8)
    return res


def plot_mask(img, mask, colors=None):
    """Outline of every class plane drawn over ``img`` (segmenation.py:82-96).

    The reference calls ``cv2.findContours(plane, RETR_LIST,
    CHAIN_APPROX_NONE)`` and then ``cv2.polylines(img, contour, True, color,
    2)`` with the contour's POINT array, i.e. one degenerate poly-line per
    contour point.  Restated without OpenCV (pinned to OpenCV 4.13 outputs in
    tests/golden/composite.npz):
      * the contour points of all borders are the mask pixels that have a
        zero 4-neighbour (the frame counts as zero);
      * a thickness-2 line from a point to itself paints that pixel and its
        four 4-neighbours (clipped to the image);
      * classes are drawn in order, later ones over earlier ones.
    A 2-D image is first replicated to three channels (cv2.COLOR_GRAY2BGR).
    """
    img = np.asarray(img)
    if img.ndim == 2:
        img = np.repeat(img[:, :, None], 3, axis=2)
    img = img.astype(np.uint8).copy()
    for idx, plane in enumerate(np.asarray(mask).astype(np.uint8)):
        m = plane != 0
        p = np.pad(m, 1)
        border = m & ~(p[:-2, 1:-1] & p[2:, 1:-1] & p[1:-1, :-2] & p[1:-1, 2:])
        b = np.pad(border, 1)
        stamp = (b[1:-1, 1:-1] | b[:-2, 1:-1] | b[2:, 1:-1] | b[1:-1, :-2] |
                 b[1:-1, 2:])
        img[stamp] = colors[idx] if colors else (0, 255, 0)
    return img
