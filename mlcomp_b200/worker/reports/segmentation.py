"""``SegmentationReportBuilder`` - mlcomp/worker/reports/segmenation.py (sic).

Same constructor (it calls ``create_base`` itself, :53) and methods.  Only
``plot_count`` images are ever drawn, so compositing / contours / JPEG stay
on the CPU with OpenCV as in the reference.
"""
import numpy as np
import yaml

from mlcomp_b200.worker.reports.classification import resize_saving_ratio
from mlcomp_b200.worker.store import (Report, ReportImg, ReportSeries,
                                      ReportTasks, now)


def _host(x):
    return x.detach().cpu().numpy() if hasattr(x, 'detach') else np.asarray(x)


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
        mask = _host(mask)
        res = np.zeros((*mask.shape[1:], 3), dtype=np.uint8)
        for i, plane in enumerate(mask):
            color = self.colors[i] if self.colors is not None else \
                (255, 255, 255)
            res += (plane[:, :, None] * np.asarray(color)).astype(np.uint8)
        return res

    def plot_mask(self, img, mask):          # segmenation.py:82-96
        import cv2
        img, mask = _host(img), _host(mask)
        if img.ndim == 2:
            img = cv2.cvtColor(img, cv2.COLOR_GRAY2BGR)
        img = np.ascontiguousarray(img.astype(np.uint8))
        for i, plane in enumerate(mask.astype(np.uint8)):
            contours, _ = cv2.findContours(plane, cv2.RETR_LIST,
                                           cv2.CHAIN_APPROX_NONE)
            color = self.colors[i] if self.colors else (0, 255, 0)
            for contour in contours:
                cv2.polylines(img, contour, True, color, 2)
        return img

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
