"""contrib/torch/tensors.py:4-8 ``flip`` on the device, as index arithmetic
inside a kernel (no index tensor, no gather)."""
from mlcomp_b200 import ops


def flip(x, dim):
    """Reverse ``x`` along one of its last two dims (what TtaWrap.inverse
    asks for).  CUDA fp16/fp32 tensors of rank >= 2."""
    nd = x.dim()
    dim = dim % nd
    if dim < nd - 2 or nd < 2:
        raise ValueError('flip: only the last two dims are supported')
    tag = 'hflip' if dim == nd - 1 else 'vflip'
    y = x.contiguous().reshape(1, -1, x.shape[-2], x.shape[-1])
    out = ops.seg_prob([y], [[tag]], None)
    return out.to(x.dtype).reshape(x.shape)


__all__ = ['flip']
