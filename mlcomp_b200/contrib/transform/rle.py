"""contrib/transform/rle.py:4-31 on the device."""
import numpy as np
import torch

from mlcomp_b200 import ops


def mask2rle(img) -> str:
    """img: (H, W) array or tensor, 1 - mask, 0 - background.  Returns the
    run-length string (column-major, 1-based "start length" pairs)."""
    t = torch.as_tensor(np.ascontiguousarray(img) if isinstance(img, np.ndarray)
                        else img)
    if t.dim() != 2:
        raise ValueError('mask2rle expects one (H, W) mask')
    t = (t != 0).to(torch.uint8).cuda()[None, None].contiguous()
    runs, off, cnt, _ = ops.rle_encode([t], 0.5, 0, activation=None)
    return ops.rle_strings(runs, off, cnt)[0]


def masks2rle(masks):
    """Batched form: [n, C, H, W] {0,1} CUDA tensor -> list of n*C strings
    from one kernel launch."""
    runs, off, cnt, _ = ops.rle_encode([masks.to(torch.uint8).contiguous()],
                                       0.5, 0, activation=None)
    return ops.rle_strings(runs, off, cnt)


def rle2mask(mask_rle, shape):
    """rle.py:16-31: ``shape`` = (width, height); returns (height, width)
    uint8.  Decoding builds ground truth off the hot path and stays on the
    host (a difference array instead of the reference's Python loop)."""
    tok = np.asarray(mask_rle.split(), dtype=np.int64)
    total = int(shape[0]) * int(shape[1])
    delta = np.zeros(total + 1, dtype=np.int32)
    starts = tok[0::2] - 1
    ends = np.minimum(starts + tok[1::2], total)
    np.add.at(delta, starts, 1)
    np.add.at(delta, ends, -1)
    return (np.cumsum(delta[:-1]) > 0).astype(np.uint8).reshape(shape).T


def rles2masks(strings, shape):
    """Batched decoder on the device: list of RLE strings -> uint8 CUDA
    tensor [len(strings), height, width]; ``shape`` = (width, height) as in
    rle2mask."""
    runs, off, cnt = ops.rle_parse(strings)
    return ops.rle_decode(runs, off, cnt, int(shape[1]), int(shape[0]))


__all__ = ['mask2rle', 'masks2rle', 'rle2mask', 'rles2masks']
