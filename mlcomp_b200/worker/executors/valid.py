"""``Valid`` - chunked predict -> score -> plot loop.

Hooks, as used by examples/digit-recognizer/executors/valid.py:32-83:
``create_base()``, ``count()``, ``adjust_part(part)``, ``score(preds)`` ->
per-sample array, ``score_final()`` -> float, ``plot(preds, scores)``,
``plot_final(score)``.  The final score goes to ``Task.score``
(db/models/task.py:37) and, YAML-serialisable, into the result dict that
``worker/tasks.py:233-235`` stores in ``task.result``.
"""
from .equation import Equation


class Valid(Equation):
    def score(self, preds):
        raise NotImplementedError

    def score_final(self):
        raise NotImplementedError

    def plot(self, preds, scores):
        pass

    def plot_final(self, score):
        pass

    def result_extra(self):
        """More (plain Python) entries for the result dict."""
        return {}

    def work(self) -> dict:
        self.create_base()
        parts = self.generate_parts(self.count())
        for i, part in enumerate(self.tqdm(parts, desc='valid', interval=10)):
            if self.step:
                self.step.start(1, 'part', i)
            self.set_part(part)
            preds = self.solve('y')
            scores = self.score(preds)
            self.plot(preds, scores)
        if self.step:
            self.step.end(1)
        score = float(self.score_final())
        self.plot_final(score)
        if self.task is not None:
            self.task.score = score
        res = {'res': score}
        res.update(self.result_extra())
        return res


__all__ = ['Valid']
