"""The boundary claim "a DAG YAML runs unchanged", checked with the
reference's OWN user code: the class bodies of
/root/reference/examples/digit-recognizer/executors/{valid,infer}.py are
executed as they are (oracle/make_ref.py copies the three example files into
the git-ignored oracle/_ref/examples/, which travels to the GPU box; the test
is skipped only where that recipe never ran) on the bases this package provides under the reference's module names
(mlcomp_b200.install_as_mlcomp, worker/storage.py:249-306 imports executors
by module path).  Only what is absent from this image is stubbed: the
``dataset`` / ``experiment`` modules of the example (catalyst, csv files) and
the NumPy aliases ``np.float`` / ``np.int`` that NumPy 2 removed.
"""
import os
import pickle
import sys
import types

import numpy as np
import pytest
import torch

REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(
    __file__))), 'oracle', '_ref', 'examples')


def test_shim_provides_the_reference_module_names():
    import mlcomp_b200
    names = mlcomp_b200.install_as_mlcomp()
    from mlcomp.worker.executors import Executor          # noqa: F401
    from mlcomp.worker.executors.valid import Valid       # noqa: F401
    from mlcomp.worker.executors.infer import Infer       # noqa: F401
    from mlcomp.worker.executors.equation import Equation  # noqa: F401
    from mlcomp.worker.reports.classification import \
        ClassificationReportBuilder                       # noqa: F401
    from mlcomp.worker.reports.segmenation import \
        SegmentationReportBuilder                         # noqa: F401
    from mlcomp.contrib.metrics.dice import dice_numpy    # noqa: F401
    from mlcomp.contrib.transform.rle import mask2rle, rle2mask  # noqa: F401
    from mlcomp.contrib.transform.tta import TtaWrap      # noqa: F401
    from mlcomp.contrib.torch.tensors import flip         # noqa: F401
    from mlcomp.utils.torch import apply_activation, infer  # noqa: F401
    assert 'mlcomp.worker.executors.valid' in names
    import mlcomp
    assert mlcomp.__path__ == []        # nothing on disk, no side effects
    assert not os.path.exists(os.path.expanduser('~/mlcomp/configs'))


def test_parse_albu_short_dict_form():
    from mlcomp_b200.contrib.transform.tta import TtaWrap, parse_albu_short
    assert parse_albu_short('hflip') == 'hflip'
    assert parse_albu_short({'name': 'VerticalFlip', 'p': 1}) == 'vflip'
    assert parse_albu_short({'name': 'Compose', 'items': [
        {'name': 'Transpose'}, 'hflip']}) == ['transpose', 'hflip']
    with pytest.raises(Exception, match='Unknwon augmentation'):
        parse_albu_short('rot90')
    with pytest.raises(AssertionError):
        parse_albu_short(['hflip'])
    with pytest.raises(Exception, match='no test-time inverse'):
        parse_albu_short({'name': 'RandomCrop'})
    assert TtaWrap([1, 2], ['hflip', 'vflip']).tags == ['hflip', 'vflip']


def test_infer_best_callback_files(tmp_path):
    """contrib/catalyst/callbacks/inference.py:10-49: same hooks, same
    ``{out_prefix}/{loader}.{key}.npy`` naming, best_only honoured."""
    from mlcomp_b200.contrib.catalyst.callbacks.inference import \
        InferBestCallback

    class Metrics:
        is_best = True

    class State:
        out_dir = str(tmp_path)
        out_prefix = 'pred'
        loader_name = 'valid'
        metrics = Metrics()
        output = None

    cb = InferBestCallback()
    st = State()
    cb.on_stage_start(st)
    cb.on_loader_start(st)
    parts = [torch.arange(6.).reshape(2, 3) + k for k in range(3)]
    for p in parts:
        st.output = {'logits': p, 'aux': p * 2}
        cb.on_batch_end(st)
    cb.on_loader_end(st)
    got = np.load(tmp_path / 'pred' / 'valid.logits.npy')
    assert np.array_equal(got, torch.cat(parts).numpy())
    assert np.array_equal(np.load(tmp_path / 'pred' / 'valid.aux.npy'),
                          got * 2)
    assert torch.equal(cb.sources['logits'], torch.cat(parts))
    # not the best epoch: nothing is written
    st.metrics.is_best = False
    cb2 = InferBestCallback(out_dir=str(tmp_path), out_prefix='p2',
                            best_only=True)
    cb2.on_stage_start(st)
    cb2.on_loader_start(st)
    st.output = {'logits': parts[0]}
    cb2.on_batch_end(st)
    cb2.on_loader_end(st)
    assert not os.path.exists(tmp_path / 'p2' / 'valid.logits.npy')


def _load_reference_class(path, name):
    """exec the reference file with the example's two local modules stubbed."""
    src = open(path).read()
    mod = types.ModuleType('ref_example_' + name)
    exec(compile(src, path, 'exec'), mod.__dict__)
    return getattr(mod, name)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.isfile(os.path.join(REF, 'valid.py')),
                    reason='oracle/_ref/examples is absent: run '
                           'python oracle/make_ref.py where /root/reference '
                           'exists')
def test_reference_valid_and_infer_mnist_run_on_our_bases(tmp_path,
                                                          monkeypatch):
    import mlcomp_b200
    mlcomp_b200.install_as_mlcomp()
    from mlcomp_b200.worker import store
    from mlcomp_b200.worker.executors import Executor
    import oracle
    if not hasattr(np, 'float'):
        monkeypatch.setattr(np, 'float', float, raising=False)
    if not hasattr(np, 'int'):
        monkeypatch.setattr(np, 'int', int, raising=False)
    monkeypatch.chdir(tmp_path)
    os.makedirs('data')
    os.makedirs('models')
    # MNIST-shaped synthetic csv files (dataset.py reads these)
    import pandas as pd
    rng = np.random.default_rng(0)
    n = 300
    y = rng.integers(0, 10, n)
    x = rng.integers(0, 255, (n, 784))
    x[np.arange(n), y * 70] = 255          # a pixel that gives the class away
    df = pd.DataFrame(x, columns=[f'pixel{i}' for i in range(784)])
    df.insert(0, 'label', y)
    df.to_csv('data/train.csv', index=False)
    pd.DataFrame({'fold': rng.integers(0, 2, n)}).to_csv('data/fold.csv',
                                                          index=False)
    df.drop(columns='label').iloc[:50].to_csv('data/test.csv', index=False)
    # the example's own dataset module, unchanged; a stub Experiment
    ds_src = open(os.path.join(REF, 'dataset.py')).read()
    dataset = types.ModuleType('dataset')
    exec(compile(ds_src, 'dataset.py', 'exec'), dataset.__dict__)
    experiment = types.ModuleType('experiment')

    class Experiment:
        @staticmethod
        def get_test_transforms():
            return lambda image: {'image': (image / 255.0 - 0.485) / 0.229}
    experiment.Experiment = Experiment
    monkeypatch.setitem(sys.modules, 'dataset', dataset)
    monkeypatch.setitem(sys.modules, 'experiment', experiment)
    # a traced model: logit c = mean of the pixels 70*c .. 70*c+9
    w = torch.zeros(10, 784)
    for c in range(10):
        w[c, 70 * c] = 1.0

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.w = torch.nn.Parameter(w)

        def forward(self, t):
            return t.reshape(t.shape[0], -1) @ self.w.t()
    torch.jit.trace(Net(), torch.zeros(2, 1, 28, 28)).save('models/net.pth')

    ValidMnist = _load_reference_class(os.path.join(REF, 'valid.py'),
                                       'ValidMnist')
    InferMnist = _load_reference_class(os.path.join(REF, 'infer.py'),
                                       'InferMnist')
    assert Executor.is_registered('valid_mnist') and \
        Executor.is_registered('infer_mnist')
    # the executor mappings of examples/digit-recognizer/all.yml:22-37
    config = {'executors': {
        'valid': {'type': 'valid_mnist', 'name': 'valid', 'plot_count': 4,
                  'y': "torch(x, file='net.pth', batch_size=64)",
                  'part_size': 100},
        'infer': {'type': 'infer_mnist', 'name': 'infer', 'test': True,
                  'prepare_submit': True, 'model_name': 'net',
                  'suffix': 'test', 'plot_count': 2,
                  'y': "torch(x, file='net.pth', batch_size=64)",
                  'part_size': 20}}}
    sess = store.MemorySession()
    task = store.Task(name='valid')
    task.id = 1
    ex = Executor.from_config(executor='valid', config=config,
                              additional_info={}, session=sess)
    assert type(ex).__name__ == 'ValidMnist'
    res = ex(task=task, task_provider=sess, dag=store.Dag(name='d'))
    # oracle: the same model on the CPU, examples/.../valid.py:45-52
    fold = pd.read_csv('data/fold.csv')['fold'].values
    keep = fold == 0
    xs = ((x[keep].reshape(-1, 1, 28, 28).astype(np.float32) / 255.0
           - 0.485) / 0.229)
    logits = xs.reshape(len(xs), -1) @ w.numpy().T
    want = oracle.classify_score_final(oracle.classify_scores(logits, y[keep]))
    assert abs(res['res'] - want) < 1e-12 and task.score == res['res']
    rows = sess.rows(store.ReportImg)
    cm_rows = [r for r in rows if r.group.endswith('_confusion')]
    assert cm_rows, 'confusion matrix row is missing'
    got_cm = sum(pickle.loads(r.img)['data'] for r in cm_rows)
    assert np.array_equal(got_cm, oracle.confusion(y[keep], logits))
    # Infer: pickle of the predictions + Kaggle csv, infer.py:62-79
    task2 = store.Task(name='infer')
    task2.id = 2
    ex2 = Executor.from_config(executor='infer', config=config,
                               additional_info={}, session=sess)
    ex2(task=task2, task_provider=sess, dag=store.Dag(name='d'))
    sub = pd.read_csv('data/submissions/net_test.csv')
    xt = ((x[:50].reshape(-1, 784).astype(np.float32) / 255.0 - 0.485)
          / 0.229) @ w.numpy().T
    assert sub['ImageId'].tolist() == list(range(1, 51))
    assert sub['Label'].tolist() == xt.argmax(axis=1).tolist()
    pred = pickle.load(open('data/pred/net_test.p', 'rb'))
    np.testing.assert_allclose(pred, xt, rtol=1e-5, atol=1e-5)
