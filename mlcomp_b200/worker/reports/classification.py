"""``ClassificationReportBuilder`` - mlcomp/worker/reports/classification.py
with the numeric lines (arg-max :111, confusion matrix :133-138) on the GPU.

Same constructor and methods.  ``session`` is anything offering the provider
calls used below (``mlcomp_b200.worker.store.MemorySession`` in this repo, the
reference's providers in MLComp).  Rows have the reference's column names.
"""
import pickle

import numpy as np
import torch
import yaml

from mlcomp_b200 import ops
from mlcomp_b200.worker.store import (Report, ReportImg, ReportSeries,
                                      ReportTasks, now)


def fit_within(img, limit):
    """Shrink ``img`` (H, W[, ch]) so that it fits ``limit = (max_h, max_w)``,
    one axis after the other, keeping the aspect ratio and truncating the
    scaled length - the two-step rule of mlcomp/utils/img.py:7-16 (a falsy
    limit, or a falsy entry of it, leaves that axis alone)."""
    if not limit:
        return img
    import cv2
    for axis in (0, 1):
        cap = limit[axis]
        if not cap or img.shape[axis] <= cap:
            continue
        other = int(cap / img.shape[axis] * img.shape[1 - axis])
        h, w = (cap, other) if axis == 0 else (other, cap)
        img = cv2.resize(img, (w, h))          # cv2 takes (width, height)
    return img


resize_saving_ratio = fit_within       # the reference's name for it


def _device_preds(preds):
    t = torch.as_tensor(preds)
    if t.dtype not in (torch.float16, torch.float32):
        t = t.float()
    return t.cuda().contiguous()


class ClassificationReportBuilder:
    def __init__(self, session, task, layout: str, part: str = 'valid',
                 name: str = 'img_classify', max_img_size=None,
                 main_metric: str = 'accuracy', plot_count: int = 0):
        self.session = session
        self.task = task
        self.part = part
        self.name = name or 'img_classify'
        self.max_img_size = max_img_size
        self.main_metric = main_metric
        self.plot_count = plot_count
        self.project = session.project(task.id).id
        self.layout = session.by_name(layout)
        self.layout_dict = yaml.safe_load(self.layout.content)

    def create_base(self):                   # classification.py:55-69
        report = Report(config=yaml.safe_dump(self.layout_dict), time=now(),
                        layout=self.layout.name, project=self.project,
                        name=self.name)
        self.session.add(report)
        self.session.add(ReportTasks(report=report.id, task=self.task.id))
        self.task.report = report.id
        self.session.update()

    def process_scores(self, scores):        # classification.py:71-85
        for key, item in self.layout_dict['items'].items():
            item['name'] = key
            if item['type'] == 'series' and item['key'] in scores:
                self.session.add(ReportSeries(
                    name=item['name'], value=float(scores[item['key']]),
                    epoch=0, time=now(), task=self.task.id, part='valid',
                    stage='stage1'))

    def confusion(self, preds, targets):
        """arg-max over classes + confusion matrix of one chunk, int64
        [C, C] (rows = truth), and the arg-max itself - one kernel."""
        p = _device_preds(preds)
        y = torch.as_tensor(np.asarray(targets) if not torch.is_tensor(targets)
                            else targets).to(torch.int64).cuda()
        cm, _, am = ops.confusion(p, y, want_argmax=True)
        return cm.cpu().numpy(), am.cpu().numpy()

    def process_pred(self, imgs, preds, targets=None, attrs=None,
                     scores=None):           # classification.py:87-152
        import cv2
        for key, item in self.layout_dict['items'].items():
            item['name'] = key
            if item['type'] != 'img_classify':
                continue
            matrix = y_pred_all = None
            if targets is not None:
                matrix, y_pred_all = self.confusion(preds, targets)
            report_imgs = []
            dag = self.session.by_id(self.task.dag)
            for i in range(len(imgs)):
                if self.plot_count <= 0:
                    break
                img = resize_saving_ratio(imgs[i], self.max_img_size)
                attr = attrs[i] if attrs else {}
                y = score = None
                if targets is not None:
                    y = int(targets[i])
                    score = float(scores[self.main_metric][i])
                    y_pred = int(y_pred_all[i])
                else:
                    y_pred = int(torch.as_tensor(preds[i]).argmax())
                _, buffer = cv2.imencode('.jpg', img)
                row = ReportImg(group=item['name'], epoch=0,
                                task=self.task.id, img=buffer,
                                dag=self.task.dag, part=self.part,
                                project=self.project, y_pred=y_pred, y=y,
                                score=score, **attr)
                report_imgs.append(row)
                if dag is not None:
                    dag.img_size = (dag.img_size or 0) + row.size
            self.session.commit()
            self.session.bulk_save_objects(report_imgs)
            if targets is not None and item.get('confusion_matrix'):
                self.session.add(ReportImg(
                    group=item['name'] + '_confusion', epoch=0,
                    task=self.task.id, img=pickle.dumps({'data': matrix}),
                    project=self.project, dag=self.task.dag, part=self.part))
            self.plot_count -= 1


__all__ = ['ClassificationReportBuilder']
