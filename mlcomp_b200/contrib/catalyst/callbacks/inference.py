"""``InferBestCallback`` - mlcomp/contrib/catalyst/callbacks/inference.py:10-49
without catalyst: the producer of the ``{out_prefix}/{loader}.{key}.npy``
files that ``Equation.load`` / the ``logits:`` entries of the segmentation
executors consume.

Same hooks and the same file naming.  The difference is where the batches
live until the loader ends: the reference copies every batch to the host in
``on_batch_end`` (``value.detach().cpu().numpy()``, :32-36); here they stay on
the device and are concatenated and copied once in ``on_loader_end`` - or not
at all: ``sources`` hands the device tensors straight to the scoring kernels.

``state`` is duck-typed (catalyst's ``State`` is not required): attributes
``output`` (dict of tensors), ``loader_name``, ``metrics.is_best`` and
optionally ``out_dir`` / ``out_prefix``.
"""
import os
from collections import defaultdict

import numpy as np
import torch


class InferBestCallback:
    def __init__(self, out_dir=None, out_prefix=None, best_only=False):
        self.out_dir = out_dir
        self.out_prefix = out_prefix
        self.predictions = defaultdict(lambda: [])
        self.best_only = best_only
        self._keys_from_state = ['out_dir', 'out_prefix']
        self.sources = {}        # key -> device tensor of the last loader

    def on_stage_start(self, state):                 # inference.py:18-27
        for key in self._keys_from_state:
            value = getattr(state, key, None)
            if value is not None:
                setattr(self, key, value)
        if self.out_dir is not None:
            self.out_prefix = str(self.out_dir) + '/' + str(self.out_prefix)
        if self.out_prefix is not None:
            os.makedirs(os.path.dirname(self.out_prefix), exist_ok=True)

    def on_loader_start(self, state):                # inference.py:29-30
        self.predictions = defaultdict(lambda: [])

    def on_batch_end(self, state):                   # inference.py:32-36
        for key, value in state.output.items():
            self.predictions[key].append(value.detach())   # stays on device

    def on_loader_end(self, state):                  # inference.py:38-49
        if self.best_only and not state.metrics.is_best:
            return
        self.sources = {key: torch.cat(value, dim=0)
                        for key, value in self.predictions.items()}
        self.predictions = {key: value.cpu().numpy()
                            for key, value in self.sources.items()}
        if self.out_prefix is not None:
            os.makedirs(self.out_prefix, exist_ok=True)
            for key, value in self.predictions.items():
                suffix = '.'.join([state.loader_name, key])
                np.save(f'{self.out_prefix}/{suffix}.npy', value)


__all__ = ['InferBestCallback']
