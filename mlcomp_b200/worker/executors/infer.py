"""``Infer`` - chunked predict -> save -> submit -> plot loop.

Hooks, as used by examples/digit-recognizer/executors/infer.py:43-98:
``create_base()``, ``count()``, ``adjust_part(part)``, ``save(preds,
folder)``, ``save_final(folder)``, ``submit(preds)``, ``submit_final(folder)``,
``plot(preds)``.  ``submit_final`` writes
``{folder}/{model_name}_{suffix}.csv``, the file the ``Submit`` executor reads
(worker/executors/kaggle.py:100-102).
"""
import os

from .equation import Equation


class Infer(Equation):
    pred_folder = 'data/pred'
    submit_folder = 'data/submissions'

    def save(self, preds, folder: str):
        pass

    def save_final(self, folder: str):
        pass

    def submit(self, preds):
        pass

    def submit_final(self, folder: str):
        pass

    def plot(self, preds):
        pass

    def work(self) -> dict:
        self.create_base()
        os.makedirs(self.pred_folder, exist_ok=True)
        if self.prepare_submit:
            os.makedirs(self.submit_folder, exist_ok=True)
        parts = self.generate_parts(self.count())
        for i, part in enumerate(self.tqdm(parts, desc='infer', interval=10)):
            if self.step:
                self.step.start(1, 'part', i)
            self.set_part(part)
            preds = self.solve('y')
            self.save(preds, self.pred_folder)
            if self.prepare_submit:
                self.submit(preds)
            self.plot(preds)
        if self.step:
            self.step.end(1)
        self.save_final(self.pred_folder)
        if self.prepare_submit:
            self.submit_final(self.submit_folder)
        return {}


__all__ = ['Infer']
