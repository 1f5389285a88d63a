"""``SegmentationReportBuilder`` - mlcomp/worker/reports/segmenation.py (sic).

Same constructor (it calls ``create_base`` itself, :53) and methods.  The
colour composite and the class outlines are CUDA kernels (csrc/composite.cu,
pixel-exact against the reference's NumPy / OpenCV 4.13 results); resizing,
stacking and JPEG encoding of the at most ``plot_count`` tiles stay OpenCV
host code as in the reference.
"""
import numpy as np
import torch
import yaml

from mlcomp_b200 import ops

from mlcomp_b200.worker.reports.classification import resize_saving_ratio
from mlcomp_b200.worker.store import (Report, ReportImg, ReportSeries,
                                      ReportTasks, now)


def _host(x):
    return x.detach().cpu().numpy() if hasattr(x, 'detach') else np.asarray(x)


def _DEV():
    return torch.device('cuda', torch.cuda.current_device())


def _device_u8(mask):
    """mask.astype(np.uint8) of the reference (:86), as a device tensor."""
    if not hasattr(mask, 'detach'):
        mask = torch.from_numpy(np.ascontiguousarray(mask))
    return mask.to(_DEV()).to(torch.uint8)


class SegmentationReportBuilder:
    def __init__(self, session, task, layout: str, part: str = 'valid',
                 name: str = 'img_segment', max_img_size=None,
                 stack_type: str = 'vertical', main_metric: str = 'dice',
                 plot_count: int = 0, colors=None):
        self.session = session
        self.task = task
        self.part = part
        self.name = name or 'img_segment'
        self.max_img_size = max_img_size
        self.stack_type = stack_type
        self.main_metric = main_metric
        self.colors = colors
        self.plot_count = plot_count
        self.project = session.project(task.id).id
        self.layout = session.by_name(layout)
        self.layout_dict = yaml.safe_load(self.layout.content)
        self.create_base()

    def create_base(self):                   # segmenation.py:55-69
        report = Report(config=yaml.safe_dump(self.layout_dict), time=now(),
                        layout=self.layout.name, project=self.project,
                        name=self.name)
        self.session.add(report)
        self.session.add(ReportTasks(report=report.id, task=self.task.id))
        self.task.report = report.id
        self.session.update()

    def encode_pred(self, mask):             # segmenation.py:71-80
        """(C, H, W) mask -> (H, W, 3) uint8 composite, on the device
        (mlc_mask_composite); returned as numpy for the stacking / JPEG
        steps, which stay OpenCV host code as in the reference."""
        m = _device_u8(mask)
        return ops.mask_composite(m[None], self.colors)[0].cpu().numpy()

    def plot_mask(self, img, mask):          # segmenation.py:82-96
        """Outlines of every class over the image (mlc_mask_outline)."""
        img = _host(img)
        if img.ndim == 2:                    # cv2.COLOR_GRAY2BGR
            img = np.repeat(img[:, :, None], 3, axis=2)
        dev_img = torch.from_numpy(np.ascontiguousarray(
            img.astype(np.uint8))).to(_DEV())
        out = ops.mask_outline(dev_img[None], _device_u8(mask)[None],
                               self.colors)
        return out[0].cpu().numpy()

    def process_scores(self, scores):        # segmenation.py:98-112
        for key, item in self.layout_dict['items'].items():
            item['name'] = key
            if item['type'] == 'series' and item['key'] in scores:
                self.session.add(ReportSeries(
                    name=item['name'], value=scores[item['key']], epoch=0,
                    time=now(), task=self.task.id, part='valid',
                    stage='stage1'))

    def process_pred(self, imgs, preds: dict, targets=None, attrs=None,
                     scores=None):           # segmenation.py:114-170
        import cv2
        for key, item in self.layout_dict['items'].items():
            item['name'] = key
            if item['type'] != 'img_segment':
                continue
            rows = []
            dag = self.session.by_id(self.task.dag)
            for i in range(len(imgs)):
                if self.plot_count <= 0:
                    break
                img = self.plot_mask(imgs[i], targets[i]) \
                    if targets is not None else _host(imgs[i])
                tiles = [img] + [self.encode_pred(v[i]) for v in preds.values()]
                tiles = [resize_saving_ratio(t, self.max_img_size)
                         for t in tiles]
                img = np.hstack(tiles) if self.stack_type == 'horizontal' \
                    else np.vstack(tiles)
                attr = attrs[i] if attrs else {}
                score = None
                if targets is not None:
                    score = float(scores[self.main_metric][i])
                _, buffer = cv2.imencode('.jpg', img)
                row = ReportImg(group=item['name'], epoch=0,
                                task=self.task.id, img=buffer,
                                dag=self.task.dag, part=self.part,
                                project=self.project, score=score, **attr)
                self.plot_count -= 1
                rows.append(row)
                if dag is not None:
                    dag.img_size = (dag.img_size or 0) + row.size
            self.session.commit()
            self.session.bulk_save_objects(rows)


__all__ = ['SegmentationReportBuilder']
