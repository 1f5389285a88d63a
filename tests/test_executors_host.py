"""Host-side plugin logic (no GPU): executor registry and construction
(worker/executors/base/executor.py:141-177), the reconstructed Valid / Infer /
Equation contract driven by a NumPy subclass shaped like
examples/digit-recognizer/executors/valid.py, and the row stand-ins."""
import os
import pickle

import numpy as np
import pytest
import yaml

from mlcomp_b200.worker.executors import (Equation, Executor, Infer, Valid,
                                          to_snake)
from mlcomp_b200.worker import store


@Executor.register
class ValidToy(Valid):
    """ValidMnist (valid.py:14-83) with arrays instead of the csv dataset."""

    def __init__(self, **kwargs):
        super().__init__(cache_names=['y'], layout='img_classify', **kwargs)
        rng = np.random.default_rng(0)
        self.all_y = rng.integers(0, 10, 1000)
        self.all_logits = rng.normal(0, 1, (1000, 10))
        self.all_logits[np.arange(1000), self.all_y] += 2
        self.scores = []
        self.calls = 0

    def functions(self):
        fns = super().functions()
        fns['model'] = self.model
        return fns

    def model(self, x, scale=1.0):
        self.calls += 1
        return x * scale

    def count(self):
        return len(self.all_y)

    def adjust_part(self, part):
        self.x = self.all_logits[part[0]:part[1]]
        self.y_true = self.all_y[part[0]:part[1]]

    def score(self, preds):
        res = (preds.argmax(axis=1) == self.y_true).astype(float)
        self.scores.extend(res)
        return res

    def score_final(self):
        return np.mean(self.scores)


def test_registry_names_and_from_config():
    assert to_snake('ValidToy') == 'valid_toy'
    for name in ('ValidToy', 'validtoy', 'valid_toy'):
        assert Executor.is_registered(name)
    config = {'executors': {'valid': {'type': 'valid_toy', 'gpu': 1,
                                      'y': 'model(x, scale=2.0)',
                                      'part_size': 300, 'depends': 'train',
                                      'file': 'data/train.csv'}}}
    ex = Executor.from_config(executor='valid', config=config,
                              additional_info=None, session='S', logger=None,
                              logger_db=None)
    assert isinstance(ex, ValidToy) and ex.session == 'S'
    assert ex.equations == {'y': 'model(x, scale=2.0)'}
    assert ex.file == 'data/train.csv'          # plain strings stay strings
    assert ex.gpu == 1 and ex.depends == 'train'  # YAML keys -> attributes
    with pytest.raises(ModuleNotFoundError, match='has not been found'):
        Executor.from_config(executor='nope', config=config,
                             additional_info=None)


def test_valid_work_loop_parts_cache_and_result():
    ex = ValidToy(y='model(x, scale=2.0)', part_size=300, name='v')
    task = store.Task(name='t')
    task.id = 7
    sess = store.MemorySession()
    res = ex(task=task, task_provider=sess, dag=store.Dag(name='d'))
    want = float((ex.all_logits.argmax(1) == ex.all_y).mean())
    assert res == {'res': want} and task.score == want
    assert ex.calls == 4                         # 1000 in parts of 300
    assert yaml.safe_load(yaml.safe_dump(res)) == res   # tasks.py:233-235
    assert task.batch_total == 4 and task.batch_index == 4
    # cache_names=['y']: a second solve in the same part is not re-run
    ex.solve('y')
    assert ex.calls == 4
    assert ex.generate_parts(10) == [(0, 10)] or ex.part_size


def test_equation_arithmetic_and_errors():
    ex = ValidToy(y1='model(x)', y2='model(x, scale=3)', y='(y1 + y2) / 2',
                  max_count=10)
    ex.set_part((0, 4))
    np.testing.assert_allclose(ex.solve('y'), ex.all_logits[:4] * 2)
    assert ex.generate_parts(ex.count()) == [(0, 10)]
    bad = ValidToy(y='nosuch(x)')
    bad.set_part((0, 1))
    with pytest.raises(NameError, match='nosuch'):
        bad.solve('y')
    with pytest.raises(KeyError):
        ex.solve('z')


@Executor.register
class InferToy(Infer):
    def __init__(self, **kwargs):
        super().__init__(cache_names=['y'], layout='img_classify', **kwargs)
        self.data = np.arange(50, dtype=np.float64).reshape(25, 2)
        self.res, self.submit_res = [], []

    def functions(self):
        return dict(super().functions(), ident=lambda v: v)

    def count(self):
        return len(self.data)

    def adjust_part(self, part):
        self.x = self.data[part[0]:part[1]]

    def save(self, preds, folder):
        self.res.extend(preds)

    def save_final(self, folder):
        pickle.dump(np.array(self.res),
                    open(f'{folder}/{self.model_name}_{self.suffix}.p', 'wb'))

    def submit(self, preds):
        self.submit_res.extend(
            [{'ImageId': len(self.submit_res) + i + 1, 'Label': int(p)}
             for i, p in enumerate(preds.argmax(axis=1))])   # infer.py:71-74

    def submit_final(self, folder):
        import pandas as pd
        pd.DataFrame(self.submit_res).to_csv(
            f'{folder}/{self.model_name}_{self.suffix}.csv', index=False)


def test_infer_work_writes_what_submit_reads(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    ex = InferToy(y='ident(x)', part_size=10, prepare_submit=True,
                  model_name='net', suffix='test', model_id=3)
    assert hasattr(ex, 'model_id') and ex.suffix == 'test'   # kaggle.py:87-101
    ex(task=store.Task(name='t'), task_provider=store.MemorySession(),
       dag=store.Dag(name='d'))
    import pandas as pd
    df = pd.read_csv('data/submissions/net_test.csv')        # kaggle.py:100-102
    assert list(df.columns) == ['ImageId', 'Label'] and len(df) == 25
    assert list(df.ImageId) == list(range(1, 26)) and (df.Label == 1).all()
    got = pickle.load(open('data/pred/net_test.p', 'rb'))
    assert np.array_equal(got, ex.data)
    # without prepare_submit nothing is submitted
    ex2 = InferToy(y='ident(x)', model_name='n2', suffix='v')
    ex2(task=store.Task(name='t'), task_provider=store.MemorySession(),
        dag=store.Dag(name='d'))
    assert not os.path.exists('data/submissions/n2_v.csv')


def test_rows_and_report_series():
    from mlcomp_b200.worker.reports import ClassificationReportBuilder
    sess = store.MemorySession()
    task = store.Task(name='t', dag=1)
    task.id = 5
    b = ClassificationReportBuilder(sess, task, 'img_classify', name=None,
                                    plot_count=2)
    assert b.name == 'img_classify'
    b.create_base()
    assert task.report == 1 and sess.rows(store.ReportTasks)[0].task == 5
    b.process_scores({'accuracy': np.float64(0.5), 'other': 1})
    (s,) = sess.rows(store.ReportSeries)
    assert (s.name, s.value, s.epoch, s.part, s.stage) == \
        ('accuracy', 0.5, 0, 'valid', 'stage1')
    with pytest.raises(TypeError):
        store.ReportImg(nonsense=1)
    with pytest.raises(KeyError):
        ClassificationReportBuilder(sess, task, 'missing-layout')


def test_rle2mask_host_decoder_matches_golden(golden_dir):
    import json
    from mlcomp_b200.contrib.transform.rle import rle2mask
    z = np.load(os.path.join(golden_dir, 'rle_masks.npz'))
    rows = json.load(open(os.path.join(golden_dir, 'rle_rows.json')))
    off = 0
    for r in rows:
        h, w = r['h'], r['w']
        m = z['mask'][off:off + h * w].reshape(h, w)
        off += h * w
        assert np.array_equal(rle2mask(r['rle'], (w, h)), m)


def test_palette_wraps_like_numpy_and_saturates_like_opencv():
    """Host side of mask_composite / mask_outline: encode_pred multiplies in
    uint8 (modulo 256), cv2 colour scalars saturate; a palette shorter than
    the class count fails like ``colors[i]`` does in the reference."""
    from mlcomp_b200.ops import _palette
    wrap, _ = _palette([(300, -1, 255), (0, 256, 511)], 2, wrap=True)
    assert wrap.tolist() == [[44, 255, 255], [0, 0, 255]]
    sat, _ = _palette([(300, -1, 255), (0, 256, 511)], 2, wrap=False)
    assert sat.tolist() == [[255, 0, 255], [0, 255, 255]]
    assert _palette(None, 3, wrap=True) == (None, None)
    with pytest.raises(IndexError):
        _palette([(1, 2, 3)], 2, wrap=True)


def test_shard_ranges_cover_the_set_once():
    from mlcomp_b200.scoring import shard_range
    for n in (0, 1, 7, 6666, 30000):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
