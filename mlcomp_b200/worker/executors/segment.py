"""Ready-made segmentation executors on the GPU path (``type: valid_segment``
/ ``type: infer_segment`` in a DAG YAML).

The reference ships no segmentation subclass of Valid / Infer; SURVEY.md 3.3
reconstructs the flow from the leaves (TtaWrap + infer -> mean -> dice_numpy
over thresholds x min-areas -> report -> mask2rle -> csv).  These two classes
are that flow with every pixel pass inside the CUDA kernels.

Inputs are either equations (``y: torch(x, file='unet.pth', ...)`` returning
``LogitSource``s) or arrays / ``.npy`` paths given as ``logits`` (one entry
per TTA variant / ensemble member, each [N, C, H, W] fp16|fp32), ``tta`` (one
tag list per entry) and ``targets`` ([N, C, H, W] one-hot).
"""
import os

import numpy as np
import torch

from mlcomp_b200 import ops, scoring
from mlcomp_b200.utils.torch import LogitSource
from .base import Executor
from .infer import Infer
from .valid import Valid


def _open(x):
    if isinstance(x, str):
        return np.load(x, mmap_mode='r')
    return x


def _to_device(x, dtype=None):
    if isinstance(x, LogitSource):
        x = x.tensor
    if not torch.is_tensor(x):
        x = torch.from_numpy(np.ascontiguousarray(x))
    if dtype is not None and x.dtype != dtype:
        x = x.to(dtype)
    elif dtype is None and x.dtype not in (torch.float16, torch.float32):
        x = x.float()
    return x.cuda(non_blocking=True).contiguous()


def _is_host_array(x):
    return isinstance(x, np.ndarray) or (torch.is_tensor(x) and not x.is_cuda)


class _SegInputs:
    """Shared plumbing: sources, tags, part slicing.

    Host-resident inputs (arrays, pinned tensors, ``.npy`` paths) are staged
    by a ``scoring.HostStager``: the part after the current one is already on
    its way to the device while the current one is scored."""

    def _init_inputs(self, logits, tta, activation):
        self.sources = [_open(s) for s in (logits or [])]
        self.tta = [list(t) for t in (tta or [[]] * len(self.sources))]
        self.activation = activation
        self._stager = None
        self._parts = []
        self._extra_host = []      # more arrays staged with the sources

    def count(self):
        if self.sources:
            return len(self.sources[0])
        return len(self.x)          # dataset attribute used by equations

    def generate_parts(self, count):
        self._parts = super().generate_parts(count)
        return self._parts

    def _staged(self):
        """Device views of the current part of sources + _extra_host, or
        None when the inputs are not plain host arrays."""
        arrays = list(self.sources) + list(self._extra_host)
        if not arrays or not all(_is_host_array(a) for a in arrays) or \
                not all(a.dtype in (np.float16, np.float32, np.uint8,
                                    torch.float16, torch.float32, torch.uint8)
                        for a in arrays):
            return None
        if self._stager is None:
            size = max(hi - lo for lo, hi in self._parts) if self._parts \
                else self.part[1] - self.part[0]
            self._stager = scoring.HostStager(arrays, 'cuda', size)
        # whatever was enqueued on the previous part's buffers so far is what
        # must finish before they are overwritten (a no-op when score() has
        # released them already)
        self._stager.release()
        dev = self._stager.fetch(*self.part)
        if self.part in self._parts:
            i = self._parts.index(self.part)
            if i + 1 < len(self._parts):
                self._stager.prefetch(*self._parts[i + 1])
        return dev

    def _release_part(self):
        if self._stager is not None:
            self._stager.release()

    def solve(self, name):
        if name == 'y' and 'y' not in self.equations:
            lo, hi = self.part
            dev = self._staged()
            self._staged_extra = None
            if dev is not None:
                self._staged_extra = dev[len(self.sources):]
                return [LogitSource(d, t, self.activation)
                        for d, t in zip(dev, self.tta)]
            return [LogitSource(_to_device(s[lo:hi]), t, self.activation)
                    for s, t in zip(self.sources, self.tta)]
        return super().solve(name)

    @staticmethod
    def _as_sources(preds):
        if isinstance(preds, LogitSource):
            preds = [preds]
        tensors = [_to_device(p) for p in preds]
        tags = [list(getattr(p, 'tags', [])) for p in preds]
        act = getattr(preds[0], 'activation', 'sigmoid')
        return tensors, tags, act


@Executor.register
class ValidSegment(_SegInputs, Valid):
    def __init__(self, logits=None, tta=None, targets=None, imgs=None,
                 bins=100, min_areas=(0,), activation='sigmoid',
                 report_threshold=0.5, colors=None, targets_rle=None,
                 argmax=False, **kwargs):
        kwargs.setdefault('layout', 'img_segment')
        super().__init__(**kwargs)
        self._init_inputs(logits, tta, activation)
        self.targets = _open(targets)
        # ground truth in its at-rest form: a csv with an ``EncodedPixels``
        # column (one row per image x class, in that order; what
        # contrib/transform/rle.py:16-31 ``rle2mask`` is fed from) or the list
        # of strings itself - parsed once, decoded per part on the device
        self.targets_rle = targets_rle
        self._gt_runs = None
        self.imgs = imgs
        self.thr = ops.threshold_grid(int(bins))
        self.min_areas = min_areas
        self.report_k = int(np.searchsorted(self.thr,
                                            np.float32(report_threshold)))
        self.report_k = min(self.report_k, len(self.thr) - 1)
        self.colors = colors
        # argmax=True: the multiclass deployment - per pixel the first maximal
        # class of the fused scores (activation 'softmax' as a rule); the
        # masks are (label == c), there is no threshold to search, the score
        # is the mean over images and classes of their dice (empty_one rule)
        self.argmax = bool(argmax)
        self._dice_rows = []
        self.search = None
        self.builder = None
        self.scores = []
        self.best = None
        # prepare_submit: also write the run-length csv of this very set at
        # the point the search picks (out-of-fold predictions).  The per-pixel
        # bin indices of every part stay on the device (1 B/px) until the
        # thresholds are known; no second pass over the logits.
        self._kept = []
        self.rows = []
        self.names = kwargs.get('names')
        self.id_column = kwargs.get('id_column', 'ImageId')
        self.rle_column = kwargs.get('rle_column', 'EncodedPixels')
        self.d2h_bytes = 0
        if self.targets is not None and self.targets_rle is None and \
                _is_host_array(self.targets):
            self._extra_host = [self.targets]

    def create_base(self):
        if self.session is not None and self.task is not None:
            from mlcomp_b200.worker.reports import SegmentationReportBuilder
            self.builder = SegmentationReportBuilder(
                session=self.session, task=self.task, layout=self.layout,
                name=self.name, plot_count=self.plot_count,
                colors=self.colors)

    def _score_argmax(self, tensors, tags, act, gt):
        out = ops.seg_argmax(tensors, tags, act, gt_onehot=gt,
                             want_label=False, want_counts=True,
                             want_onehot=self.prepare_submit)
        counts = out['counts'].cpu().numpy()             # [n, C, 3] T, P, I
        self.d2h_bytes += counts.nbytes
        t, p, i = (counts[..., k].astype(np.float64) for k in range(3))
        dice = np.where((t == 0) & (p == 0), 1.0, 2 * i / (t + p + 1e-6))
        self._dice_rows.append(dice)                     # dice.py:17-23
        if self.prepare_submit:
            self._kept.append((self.part[0], out['onehot']))
        self._release_part()
        res = dice.mean(axis=1)
        self.scores.extend(res.tolist())
        return res

    def score(self, preds):
        tensors, tags, act = self._as_sources(preds)
        n_cls = tensors[0].shape[1]
        if self.argmax:
            lo, hi = self.part
            extra = getattr(self, '_staged_extra', None)
            gt = extra[0] if extra else self._targets(lo, hi, tensors[0],
                                                      tags[0])
            return self._score_argmax(tensors, tags, act, gt)
        if self.search is None:
            self.search = scoring.SegSearch(n_cls, self.thr, self.min_areas,
                                            activation=act)
        lo, hi = self.part
        extra = getattr(self, '_staged_extra', None)
        gt = extra[0] if extra else self._targets(lo, hi, tensors[0], tags[0])
        if self.prepare_submit:
            hist, bins = self.search.add(tensors, gt, tags, keep_bins=True)
            self._kept.append((lo, hist, bins))
        else:
            hist = self.search.add(tensors, gt, tags)
        self._release_part()
        per_plane = ops.plane_dice(hist, [self.report_k] * n_cls, [0] * n_cls)
        res = per_plane.mean(dim=1).cpu().numpy()
        self.d2h_bytes += res.nbytes
        self.scores.extend(res.tolist())
        return res

    def _targets(self, lo, hi, like, tags0):
        """uint8 planes [hi-lo, C, H, W] of the part on the device."""
        if self.targets_rle is None:
            return _to_device(self.targets[lo:hi], torch.uint8)
        c = like.shape[1]
        h, w = like.shape[2:]
        if ops.compose_tta(tags0) & ops.MLC_XF_TRANSPOSE:
            h, w = w, h
        if self._gt_runs is None:
            rows = self.targets_rle
            if isinstance(rows, str):
                import pandas as pd
                rows = pd.read_csv(rows, keep_default_na=False)[
                    'EncodedPixels'].tolist()
            runs, off, cnt = ops.rle_parse([str(r) for r in rows])
            self._gt_runs = (runs, off.reshape(-1), cnt.reshape(-1))
        runs, off, cnt = self._gt_runs
        return ops.rle_decode(runs, off[lo * c:hi * c], cnt[lo * c:hi * c],
                              h, w).view(hi - lo, c, h, w)

    def score_final(self):
        if self.argmax:
            dice = np.concatenate(self._dice_rows)           # [N, C]
            per_class = dice.mean(axis=0)
            self.best = [{'class': c, 'threshold': None, 'min_area': 0,
                          'threshold_index': None, 'dice': float(d)}
                         for c, d in enumerate(per_class)]
            if self.prepare_submit:
                for lo, onehot in self._kept:
                    runs, off, cnt, _ = ops.rle_encode([onehot], 0.5, 0,
                                                       activation=None)
                    self._rows_from(lo, onehot.shape[0], onehot.shape[1],
                                    ops.rle_strings(runs, off, cnt))
                self._kept = []
                self._write_rows()
            return float(per_class.mean())
        out = self.search.result()
        self.best = out['best']
        self.d2h_bytes += out['score'].nbytes + 8 * (
            out['micro_P'].size + out['micro_I'].size + out['micro_T'].size)
        if self.prepare_submit:
            self._submit_kept()
        return out['dice']

    def _submit_kept(self):
        """Run-length strings of every kept part at the chosen point, read
        from the bin planes (mask = bin > k); planes whose mask is empty or
        under the min-area are known from the histograms and not read."""
        k_c = [b['threshold_index'] for b in self.best]
        a_c = [b['min_area'] for b in self.best]
        for lo, hist, bins in self._kept:
            runs, off, cnt, _ = ops.rle_from_bins(bins, k_c, a_c, hist=hist)
            strings = ops.rle_strings(runs, off, cnt)
            self._rows_from(lo, hist.shape[0], hist.shape[1], strings)
        self._kept = []
        self._write_rows()

    def _rows_from(self, lo, n, c, strings):
        self.d2h_bytes += sum(len(t) for t in strings) + 12 * len(strings)
        for i in range(n):
            name = self.names[lo + i] if self.names is not None else lo + i
            for j in range(c):
                self.rows.append({self.id_column: f'{name}_{j + 1}',
                                  self.rle_column: strings[i * c + j]})

    def _write_rows(self):
        folder = getattr(self, 'submit_folder', None)
        if folder:
            import pandas as pd
            os.makedirs(folder, exist_ok=True)
            path = os.path.join(folder,
                                f'{self.model_name}_{self.suffix}.csv')
            pd.DataFrame(self.rows, columns=[self.id_column,
                                             self.rle_column]) \
                .to_csv(path, index=False)

    def plot(self, preds, scores):
        if self.builder is None or self.builder.plot_count <= 0 or \
                self.imgs is None:
            return
        lo, hi = self.part
        k = min(hi - lo, self.builder.plot_count)
        tensors, tags, act = self._as_sources(preds)
        prob = ops.seg_prob([t[:k].contiguous() for t in tensors], tags, act)
        mask = (prob >= float(self.thr[self.report_k])).to(torch.uint8)
        self.builder.process_pred(
            imgs=self.imgs[lo:lo + k], preds={'pred': mask.cpu().numpy()},
            targets=self._targets(lo, lo + k, tensors[0],
                                  tags[0]).cpu().numpy(),
            scores={'dice': scores})

    def plot_final(self, score):
        if self.builder is not None:
            self.builder.process_scores({'dice': score})

    def result_extra(self):
        if self.argmax:
            return {'dice_per_class': [b['dice'] for b in self.best]}
        return {'thresholds': [b['threshold'] for b in self.best],
                'min_areas': [b['min_area'] for b in self.best],
                'dice_per_class': [b['dice'] for b in self.best]}


@Executor.register
class InferSegment(_SegInputs, Infer):
    def __init__(self, logits=None, tta=None, names=None, thresholds=None,
                 min_areas=None, activation='sigmoid', id_column='ImageId',
                 rle_column='EncodedPixels', **kwargs):
        kwargs.setdefault('layout', 'img_segment')
        super().__init__(**kwargs)
        self._init_inputs(logits, tta, activation)
        self.names = names
        self.thresholds = thresholds
        self.min_areas = min_areas
        self.id_column = id_column
        self.rle_column = rle_column
        self.rows = []

    def create_base(self):
        if self.thresholds is None and self.task_provider is not None:
            for res in self.dependent_results().values():
                if res and 'thresholds' in res:      # a Valid upstream
                    self.thresholds = res['thresholds']
                    self.min_areas = res.get('min_areas')
                    break
        if self.thresholds is None:
            self.thresholds = 0.5
        if self.min_areas is None:
            self.min_areas = 0

    def submit(self, preds):
        tensors, tags, act = self._as_sources(preds)
        runs, off, cnt, _ = ops.rle_encode(tensors, self.thresholds,
                                           self.min_areas, xforms=tags,
                                           activation=act)
        strings = ops.rle_strings(runs, off, cnt)
        lo, _ = self.part
        n, c = tensors[0].shape[:2]
        for i in range(n):
            name = self.names[lo + i] if self.names is not None else lo + i
            for j in range(c):
                self.rows.append({self.id_column: f'{name}_{j + 1}',
                                  self.rle_column: strings[i * c + j]})

    def submit_final(self, folder):
        import pandas as pd
        path = os.path.join(folder, f'{self.model_name}_{self.suffix}.csv')
        pd.DataFrame(self.rows, columns=[self.id_column, self.rle_column]) \
            .to_csv(path, index=False)
        return path


__all__ = ['ValidSegment', 'InferSegment']
