"""Classification executors on the GPU path (``type: valid_classify`` /
``type: infer_classify``): the flow of
examples/digit-recognizer/executors/{valid,infer}.py with arg-max, per-sample
hits and the confusion matrix computed by mlc_confusion."""
import os
import pickle

import numpy as np
import torch

from mlcomp_b200 import scoring
from mlcomp_b200.utils.torch import LogitSource
from .base import Executor
from .infer import Infer
from .segment import _open, _to_device
from .valid import Valid


class _ClsInputs:
    def _init_inputs(self, logits):
        self.logits = _open(logits)

    def count(self):
        if self.logits is not None:
            return len(self.logits)
        return len(self.x)

    def solve(self, name):
        if name == 'y' and 'y' not in self.equations:
            lo, hi = self.part
            return LogitSource(_to_device(self.logits[lo:hi]))
        return super().solve(name)


@Executor.register
class ValidClassify(_ClsInputs, Valid):
    def __init__(self, logits=None, targets=None, imgs=None, **kwargs):
        kwargs.setdefault('layout', 'img_classify')
        super().__init__(**kwargs)
        self._init_inputs(logits)
        self.targets = _open(targets)
        self.imgs = imgs
        self.scorer = None
        self.builder = None
        self.confusion = None

    def create_base(self):
        if self.session is not None and self.task is not None:
            from mlcomp_b200.worker.reports import ClassificationReportBuilder
            self.builder = ClassificationReportBuilder(
                session=self.session, task=self.task, layout=self.layout,
                name=self.name, plot_count=self.plot_count)
            self.builder.create_base()

    def score(self, preds):
        p = _to_device(preds)
        lo, hi = self.part
        y = torch.as_tensor(np.asarray(self.targets[lo:hi])) \
            .to(torch.int64).cuda()
        if self.scorer is None:
            self.scorer = scoring.ClassifyScore(p.shape[1])
        am = self.scorer.add(p, y, want_argmax=True)
        return (am.to(torch.int64) == y).double().cpu().numpy()  # valid.py:47

    def score_final(self):
        res = self.scorer.result()
        self.confusion = res['confusion']
        return res['accuracy']                                   # valid.py:52

    def plot(self, preds, scores):
        if self.builder is None:
            return
        lo, hi = self.part
        imgs = self.imgs[lo:hi] if self.imgs is not None else []
        self.builder.process_pred(imgs=imgs, preds=_to_device(preds),
                                  targets=np.asarray(self.targets[lo:hi]),
                                  scores={'accuracy': scores})

    def plot_final(self, score):
        if self.builder is not None:
            self.builder.process_scores({'accuracy': score})

    def result_extra(self):
        return {'confusion': self.confusion.tolist()}


@Executor.register
class InferClassify(_ClsInputs, Infer):
    def __init__(self, logits=None, id_column='ImageId', label_column='Label',
                 **kwargs):
        kwargs.setdefault('layout', 'img_classify')
        super().__init__(**kwargs)
        self._init_inputs(logits)
        self.id_column = id_column
        self.label_column = label_column
        self.res = []
        self.submit_res = []

    def save(self, preds, folder):                               # infer.py:62-63
        self.res.append(_to_device(preds).cpu().numpy())

    def save_final(self, folder):                                # infer.py:65-67
        with open(os.path.join(
                folder, f'{self.model_name}_{self.suffix}.p'), 'wb') as f:
            pickle.dump(np.concatenate(self.res) if self.res
                        else np.zeros((0,)), f)

    def submit(self, preds):                                     # infer.py:69-74
        p = _to_device(preds)
        dummy = torch.zeros(p.shape[0], dtype=torch.int64, device=p.device)
        am = scoring.ClassifyScore(p.shape[1]).add(p, dummy, want_argmax=True)
        base = len(self.submit_res)
        self.submit_res.extend(
            {self.id_column: base + i + 1, self.label_column: int(v)}
            for i, v in enumerate(am.cpu().numpy()))

    def submit_final(self, folder):                              # infer.py:76-79
        import pandas as pd
        path = os.path.join(folder, f'{self.model_name}_{self.suffix}.csv')
        pd.DataFrame(self.submit_res).to_csv(path, index=False)
        return path


__all__ = ['ValidClassify', 'InferClassify']
