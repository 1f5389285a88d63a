"""contrib/metrics/dice.py:4-23 with the pixel passes on the device."""
import numpy as np
import torch

from mlcomp_b200 import ops


def _plane(x, dtype):
    t = torch.as_tensor(np.ascontiguousarray(x) if isinstance(x, np.ndarray)
                        else x)
    if t.dim() != 2:
        raise ValueError('dice_numpy scores one (H, W) plane')
    return t.to(dtype).cuda()[None, None].contiguous()


def dice_numpy(targets, outputs, threshold=None, min_area=None,
               empty_one: bool = True, eps=1e-6):
    """Same signature and results as the reference: dice of one plane at one
    (threshold, min_area) point; returns ``int 1`` for an empty/empty plane
    when ``empty_one``, otherwise fp64.  ``targets`` must be {0,1}; with
    ``threshold=None`` ``outputs`` must already be a {0,1} mask.

    For a whole validation set and a whole grid use
    ``mlcomp_b200.scoring.SegSearch`` - one pass instead of one per point.
    """
    gt = (_plane(targets, torch.float32) != 0).to(torch.uint8)
    if threshold is None:
        # the reference then sums `outputs` as given (dice.py:10-11,20): for a
        # {0,1} mask that is the integer counting done here; a soft mask would
        # need numpy's float summation order and is refused rather than
        # silently binarised
        raw = _plane(outputs, torch.float32)
        if bool(((raw != 0) & (raw != 1)).any()):
            raise ValueError('dice_numpy(threshold=None) takes a {0,1} mask; '
                             'pass probabilities together with a threshold')
        prob = raw
        thr = 0.5
    else:
        prob = _plane(outputs, torch.float32)
        thr = np.float32(threshold)
    hist = ops.seg_hist([prob], gt, [thr], activation=None)
    t, p, i = (int(v.item()) for v in (
        hist[0, 0, 1].sum(), hist[0, 0, :, 1].sum(), hist[0, 0, 1, 1]))
    if min_area and p < min_area:
        p = 0
    if empty_one and t == 0 and p == 0:
        return 1
    d = ops.plane_dice(hist, [0], [int(min_area or 0)], empty_one, eps)
    return np.float64(d.item())


__all__ = ['dice_numpy']
