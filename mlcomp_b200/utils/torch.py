"""GPU-resident counterparts of mlcomp/utils/torch.py:11-70.

The reference applies the activation, un-flips, then copies every probability
to the host (``.detach().cpu().numpy()``, torch.py:29,43).  Here ``infer``
keeps the raw logits on the device and only *records* the activation and the
test-time transforms in a ``LogitSource``; the scoring kernels apply both
while they read (mlc_seg_hist / mlc_rle_encode / mlc_seg_prob).
"""
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from mlcomp_b200 import ops


@dataclass
class LogitSource:
    """Raw model output of one TTA variant / ensemble member."""
    tensor: torch.Tensor                 # [n, C, Hs, Ws] or [n, C], on device
    tags: List[str] = field(default_factory=list)    # TtaWrap.tfms
    activation: Optional[str] = None

    def __len__(self):
        return self.tensor.shape[0]

    def __getitem__(self, sl):
        if isinstance(sl, int):
            return self._host()[sl]
        return LogitSource(self.tensor[sl], self.tags, self.activation)

    def prob(self):
        """Materialise what the reference's ``_infer`` returns."""
        return fuse([self])

    # -- array protocol: user code written against the reference receives
    # ``preds`` as the NumPy array ``infer`` returned (utils/torch.py:29) and
    # calls ``preds.argmax(axis=1)``, iterates over rows, ``np.array(...)``s
    # them (examples/digit-recognizer/executors/valid.py:45-76,
    # infer.py:62-74).  The same calls work on a LogitSource; the arg-max is
    # taken on the device (mlc_confusion), everything else materialises the
    # probabilities once.
    @property
    def shape(self):
        return tuple(self.tensor.shape)

    def _host(self):
        if getattr(self, '_np', None) is None:
            object.__setattr__(self, '_np', self.prob().cpu().numpy())
        return self._np

    def __array__(self, dtype=None, copy=None):
        a = self._host()
        return a.astype(dtype) if dtype is not None else a

    def __iter__(self):
        return iter(self._host())

    def argmax(self, axis=None):
        import numpy as np
        if self.tensor.dim() == 2 and axis in (1, -1):
            p = self.prob() if self.activation else self.tensor
            p = p.contiguous()
            if p.dtype not in (torch.float16, torch.float32):
                p = p.float()
            dummy = torch.zeros(p.shape[0], dtype=torch.int64,
                                device=p.device)
            am = ops.confusion(p, dummy, want_argmax=True)[2]
            return am.cpu().numpy().astype(np.int64)
        return np.argmax(self._host(), axis=axis)


def fuse(sources):
    """Mean of activated, un-transformed sources as fp32 [n, C, H, W]."""
    act = sources[0].activation
    ts = [s.tensor.contiguous() for s in sources]
    if ts[0].dim() == 2:                           # classification logits
        if act == 'softmax':
            # over the class dim of [n, C], as apply_activation does it
            # (utils/torch.py:16-17); mean as everywhere: fp32, left to
            # right, one divide by S
            acc = None
            for t in ts:
                p = torch.softmax(t.float(), 1)
                acc = p if acc is None else acc + p
            return acc / float(len(ts))
        ts = [t[None, None] for t in ts]
        return ops.seg_prob(ts, None, act)[0, 0]
    return ops.seg_prob(ts, [s.tags for s in sources], act)


def apply_activation(x, activation):
    """utils/torch.py:11-18 on the device: falsy -> x, 'sigmoid' through the
    fused kernel, 'softmax' over dim 1 (4-D segmentation logits through the
    same kernel, 2-D classification logits through torch), anything else
    raises the reference's error."""
    if not activation:
        return x
    if activation == 'sigmoid':
        if not x.is_cuda:
            raise TypeError('x must be a CUDA tensor (mlcomp_b200 has no CPU '
                            'path)')
        shape = x.shape
        y = x.contiguous()
        y = y.reshape(1, 1, -1, shape[-1]) if y.dim() != 4 else y
        # torch.sigmoid keeps the dtype: an fp16 tensor is evaluated in
        # fp32 and rounded back, which is what the cast does
        return ops.seg_prob([y], None, 'sigmoid').reshape(shape).to(x.dtype)
    if activation == 'softmax':
        if x.is_cuda and x.dim() == 4 and \
                x.dtype in (torch.float16, torch.float32):
            return ops.seg_prob([x.contiguous()], None, 'softmax')
        return torch.softmax(x, 1)       # [n, C] classification logits
    raise Exception(f'unknown activation = {activation}')


def _tags_of(dataset):
    from mlcomp_b200.contrib.transform.tta import TtaWrap
    return list(dataset.tags) if isinstance(dataset, TtaWrap) else []


def _forward(model, loader, activation, tags):
    for batch in loader:
        features = batch['features'].cuda(non_blocking=True)
        with torch.no_grad():
            logits = model(features)
        yield LogitSource(logits.detach(), tags, activation), batch


def infer(x, file: str, batch_size: int = 1, batch_mode: bool = False,
          activation=None, num_workers: int = 1):
    """utils/torch.py:51-70 with the result left on the device.

    Returns a ``LogitSource`` over the whole dataset, or - ``batch_mode`` - a
    generator of ``{'prob': LogitSource, 'count': n, **batch}``."""
    from torch.jit import load
    from torch.utils.data import DataLoader
    if activation not in (None, '', 'sigmoid', 'softmax'):
        raise Exception(f'unknown activation = {activation}')
    loader = DataLoader(x, batch_size=batch_size, shuffle=False,
                        num_workers=num_workers, pin_memory=True)
    model = load(file).cuda()
    tags = _tags_of(x)
    if batch_mode:
        return ({'prob': src, 'count': len(src), **batch}
                for src, batch in _forward(model, loader, activation, tags))
    parts = [src.tensor for src, _ in _forward(model, loader, activation, tags)]
    return LogitSource(torch.cat(parts), tags, activation)


__all__ = ['infer', 'apply_activation', 'LogitSource', 'fuse']
