"""GPU: the reference-facing function / plugin layer, against the committed
golden fixtures (outputs of the reference's own functions) and the oracle."""
import json
import os
import pickle

import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def test_dice_numpy_dropin_reproduces_reference_golden(golden_dir):
    """Every golden dice_numpy point (reference output, make_golden.py)
    through the GPU-backed drop-in: bit-equal, int-1 rule included."""
    from mlcomp_b200.contrib.metrics.dice import dice_numpy
    z = np.load(os.path.join(golden_dir, 'dice_planes.npz'))
    rows = json.load(open(os.path.join(golden_dir, 'dice_rows.json')))
    shapes = {r['case']: (r['h'], r['w']) for r in rows}
    planes, off = {}, 0
    for case in sorted(shapes):
        h, w = shapes[case]
        planes[case] = (z['gt'][off:off + h * w].reshape(h, w),
                        z['prob'][off:off + h * w].reshape(h, w))
        off += h * w
    thr = z['thr']
    rows = [r for i, r in enumerate(rows) if i % 3 == 0 or r['k'] < 0]
    for r in rows:
        gt, prob = planes[r['case']]
        if r['k'] < 0:
            got = dice_numpy(gt, (prob >= np.float32(0.5)).astype(np.uint8))
        else:
            got = dice_numpy(gt, prob, threshold=np.float32(thr[r['k']]),
                             min_area=r['min_area'], empty_one=r['empty_one'])
        assert isinstance(got, int) == r['is_int'], r
        assert float(got) == r['value'], r


def test_mask2rle_dropin_reproduces_reference_golden(golden_dir):
    from mlcomp_b200.contrib.transform.rle import (mask2rle, masks2rle,
                                                   rle2mask)
    z = np.load(os.path.join(golden_dir, 'rle_masks.npz'))
    rows = json.load(open(os.path.join(golden_dir, 'rle_rows.json')))
    off = 0
    for r in rows:
        h, w = r['h'], r['w']
        m = z['mask'][off:off + h * w].reshape(h, w)
        off += h * w
        assert mask2rle(m) == r['rle']
        assert mask2rle(torch.from_numpy(m).to(DEV)) == r['rle']
        assert np.array_equal(rle2mask(mask2rle(m), (w, h)), m)
    rng = np.random.default_rng(2)
    batch = (rng.random((3, 2, 40, 56)) < 0.4).astype(np.uint8)
    got = masks2rle(torch.from_numpy(batch).to(DEV))
    want = [oracle.rle_encode(batch[i, j]) for i in range(3) for j in range(2)]
    assert got == want


def test_rle_decode_on_device(golden_dir):
    """mlc_rle_decode == rle2mask: the reference's golden strings, encoder
    round trips on odd shapes, clipping, a plane too large for shared
    memory."""
    from mlcomp_b200 import ops
    from mlcomp_b200.contrib.transform.rle import rles2masks
    z = np.load(os.path.join(golden_dir, 'rle_masks.npz'))
    rows = json.load(open(os.path.join(golden_dir, 'rle_rows.json')))
    off = 0
    by_shape = {}
    for r in rows:
        h, w = r['h'], r['w']
        m = z['mask'][off:off + h * w].reshape(h, w)
        off += h * w
        by_shape.setdefault((h, w), []).append((r['rle'], m))
    for (h, w), items in by_shape.items():
        got = rles2masks([s for s, _ in items], (w, h)).cpu().numpy()
        for k, (s, m) in enumerate(items):
            assert np.array_equal(got[k], m), (h, w, k)
            assert np.array_equal(got[k], oracle.rle_decode(s, (w, h)))
    rng = np.random.default_rng(3)
    for h, w in ((70, 45), (256, 320), (33, 1), (1, 65), (350, 525)):
        masks = (rng.random((5, 1, h, w)) < 0.5).astype(np.uint8)
        masks[0] = 0
        masks[1] = 1
        masks[2, 0, h // 2:, :] = 1
        x = torch.from_numpy(masks).to(DEV)
        runs, o, c, _ = ops.rle_encode([x], 0.5, 0, activation=None)
        back = ops.rle_decode(runs, o.reshape(-1), c.reshape(-1), h, w)
        assert torch.equal(back, x[:, 0])
    # a run past the end is clipped like the reference's slice assignment
    runs, o, c = ops.rle_parse(['3 100', '', '1 2 2 3'])
    got = ops.rle_decode(runs, o, c, 4, 3).cpu().numpy()
    assert np.array_equal(got[0], oracle.rle_decode('3 100', (3, 4)))
    assert got[1].sum() == 0
    assert np.array_equal(got[2], oracle.rle_decode('1 2 2 3', (3, 4)))
    # 2048 x 2048: bit plane does not fit shared memory -> global path
    big = torch.zeros((1, 1, 2048, 2048), dtype=torch.uint8, device=DEV)
    big[0, 0, 100:1900, 7:2040:3] = 1
    runs, o, c, _ = ops.rle_encode([big[:, :, :1024, :1024].contiguous()], 0.5,
                                   0, activation=None)
    small = ops.rle_decode(runs, o.reshape(-1), c.reshape(-1), 1024, 1024)
    assert torch.equal(small[0], big[0, 0, :1024, :1024])
    s = oracle.rle_encode(big[0, 0].cpu().numpy())
    assert torch.equal(rles2masks([s], (2048, 2048))[0], big[0, 0])


def test_flip_tta_and_activation_dropins(golden_dir):
    from mlcomp_b200.contrib.torch.tensors import flip
    from mlcomp_b200.contrib.transform.tta import TtaWrap
    from mlcomp_b200.utils.torch import apply_activation
    z = np.load(os.path.join(golden_dir, 'flip.npz'))
    x = torch.from_numpy(z['x']).to(DEV)
    assert np.array_equal(flip(x, 3).cpu().numpy(), z['last'])
    assert np.array_equal(flip(x, 2).cpu().numpy(), z['prev'])
    assert np.array_equal(flip(x, -1).cpu().numpy(), z['last'])
    # the rest of the reference's signature: any dim, any dtype, CPU tensors
    for t in (x, x.long(), x.cpu(), x.double()):
        for d in range(-t.dim(), t.dim()):
            assert torch.equal(flip(t, d), oracle.flip_axis(t, d % t.dim()))
    from mlcomp_b200.contrib.metrics.dice import dice_numpy
    with pytest.raises(ValueError, match='takes a \\{0,1\\} mask'):
        dice_numpy(np.ones((4, 8), np.float32),
                   np.full((4, 8), 0.5, np.float32))
    w = TtaWrap(list(range(5)), tfms=['hflip', 'vflip'])
    assert len(w) == 5 and w[2] == 2
    assert np.array_equal(w.inverse(x).cpu().numpy(), z['both'])
    t = TtaWrap([], tfms=['transpose', 'hflip'])
    want = oracle.tta_inverse(x.cpu(), ['transpose', 'hflip'])
    assert torch.equal(t.inverse(x).cpu(), want)
    assert torch.equal(t.inverse(x[0]).cpu(), want[0])
    with pytest.raises(Exception, match='Unknwon augmentation'):
        TtaWrap([], tfms=['rot90'])
    lg = torch.randn(4, 3, 8, 8, device=DEV)
    assert torch.equal(apply_activation(lg, 'sigmoid'), torch.sigmoid(lg))
    assert apply_activation(lg, None) is lg
    sm4 = apply_activation(lg, 'softmax')
    assert (sm4 - torch.softmax(lg, 1)).abs().max() < 1e-6
    assert torch.equal(apply_activation(lg[:, :, 0, 0], 'softmax'),
                       torch.softmax(lg[:, :, 0, 0], 1))
    with pytest.raises(Exception, match='unknown activation = relu'):
        apply_activation(lg, 'relu')


def _seg_data(n, c, h, w, seed):
    rng = np.random.default_rng(seed)
    gt = np.zeros((n, c, h, w), np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    for i in range(n):
        for j in range(c):
            if rng.random() < 0.4:
                continue
            cy, cx = rng.integers(0, h), rng.integers(0, w)
            gt[i, j] = (((yy - cy) / (h / 4)) ** 2 +
                        ((xx - cx) / (w / 5)) ** 2 <= 1)
    base = 5 * (gt.astype(np.float32) - 0.5) + rng.normal(0, 2.0, gt.shape)
    v0 = torch.from_numpy(base.astype(np.float32)).half()
    v1 = torch.from_numpy((base + rng.normal(0, 1, gt.shape))
                          .astype(np.float32)).half()
    return gt, v0, v1


def test_valid_then_infer_segment_dag(tmp_path, monkeypatch):
    """valid_segment -> infer_segment as two DAG tasks: the search result
    travels through task.result (YAML) like worker/tasks.py:233-235 and
    executor.py:222-227, the csv is what Submit reads (kaggle.py:100-102)."""
    import yaml
    from mlcomp_b200.worker import store
    from mlcomp_b200.worker.executors import Executor
    monkeypatch.chdir(tmp_path)
    n, c, h, w = 10, 2, 48, 72
    gt, v0, v1 = _seg_data(n, c, h, w, 3)
    tags = [[], ['hflip']]
    seen = [v0, oracle.flip_axis(v1, 3).contiguous()]
    np.save('v0.npy', seen[0].numpy())
    np.save('v1.npy', seen[1].numpy())
    np.save('gt.npy', gt)
    areas = [0, 40, 400]
    config = yaml.safe_load(f"""
executors:
  valid:
    type: valid_segment
    gpu: 1
    logits: [v0.npy, v1.npy]
    tta: [[], [hflip]]
    targets: gt.npy
    bins: 100
    min_areas: {areas}
    part_size: 4
    plot_count: 3
    name: seg
  infer:
    type: infer_segment
    gpu: 1
    depends: valid
    logits: [v0.npy, v1.npy]
    tta: [[], [hflip]]
    part_size: 3
    prepare_submit: True
    model_name: unet
    suffix: test
""")
    sess = store.MemorySession()
    dag = sess.add(store.Dag(name='d', img_size=0))
    vt = sess.add(store.Task(name='valid', dag=dag.id))
    it = sess.add(store.Task(name='infer', dag=dag.id))
    vt._child_of = it.id                    # infer depends on valid

    valid = Executor.from_config(executor='valid', config=config,
                                 additional_info=None, session=sess)
    valid.imgs = np.zeros((n, h, w, 3), np.uint8)
    res = valid(task=vt, task_provider=sess, dag=dag)
    vt.result = yaml.safe_dump(res)         # tasks.py:233-235

    prob = oracle.fuse_probs(seen, tags, 'sigmoid')
    thr = oracle.threshold_grid(100)
    # the device probabilities decide; against the CPU oracle allow 1e-5
    table = oracle.dice_grid_fast(gt, prob, thr, areas)
    want = oracle.grid_scores(table)
    best = oracle.best_point(want)
    np.testing.assert_allclose(valid.search.result()['score'], want,
                               rtol=1e-5, atol=1e-7)
    assert [round(t, 6) for t in res['thresholds']] == \
        [round(float(thr[k]), 6) for k, _ in best]
    assert res['min_areas'] == [areas[a] for _, a in best]
    assert abs(res['res'] - np.mean([want[j][best[j]] for j in range(c)])) \
        < 1e-6
    assert vt.score == res['res'] and vt.report == 1
    imgs = [r for r in sess.rows(store.ReportImg) if r.group == 'img_segment']
    assert len(imgs) == 3 and imgs[0].score is not None
    assert sess.rows(store.ReportSeries)[0].name == 'dice'

    infer = Executor.from_config(executor='infer', config=config,
                                 additional_info=None, session=sess)
    infer(task=it, task_provider=sess, dag=dag)
    import pandas as pd
    df = pd.read_csv('data/submissions/unet_test.csv', keep_default_na=False)
    assert list(df.columns) == ['ImageId', 'EncodedPixels']
    assert len(df) == n * c
    from mlcomp_b200 import ops
    dprob = ops.seg_prob([s.to(DEV) for s in seen], tags).cpu().numpy()
    for i in range(n):
        for j in range(c):
            row = df.iloc[i * c + j]
            assert row.ImageId == f'{i}_{j + 1}'
            assert row.EncodedPixels == oracle.rle_submission(
                dprob[i, j], np.float32(res['thresholds'][j]),
                res['min_areas'][j])


def test_valid_classify_report_rows():
    from mlcomp_b200.worker import store
    from mlcomp_b200.worker.executors import ValidClassify, InferClassify
    rng = np.random.default_rng(4)
    n, c = 3000, 10
    y = rng.integers(0, c, n)
    logits = rng.normal(0, 1, (n, c)).astype(np.float32)
    logits[np.arange(n), y] += 2
    sess = store.MemorySession()
    dag = sess.add(store.Dag(name='d', img_size=0))
    task = sess.add(store.Task(name='v', dag=dag.id))
    ex = ValidClassify(logits=logits, targets=y, part_size=1024, plot_count=1,
                       imgs=np.zeros((n, 8, 8, 3), np.uint8), name='cls')
    ex.session = sess
    res = ex(task=task, task_provider=sess, dag=dag)
    assert res['res'] == float(oracle.classify_score_final(
        oracle.classify_scores(logits, y)))
    assert np.array_equal(np.array(res['confusion']),
                          oracle.confusion(y, logits))
    conf = [r for r in sess.rows(store.ReportImg)
            if r.group == 'img_classify_confusion']
    assert len(conf) == 3                        # one per chunk, as upstream
    first = pickle.loads(conf[0].img)['data']    # classification.py:139-150
    assert first.dtype == np.int64
    assert np.array_equal(first, oracle.confusion(y[:1024], logits[:1024]))
    plotted = [r for r in sess.rows(store.ReportImg) if r.group == 'img_classify']
    assert len(plotted) == 1024 and plotted[5].y == int(y[5])
    assert plotted[5].y_pred == int(logits[5].argmax())
    assert sess.rows(store.ReportSeries)[0].value == res['res']
    assert dag.img_size > 0


def test_infer_keeps_logits_on_device(tmp_path):
    """utils.torch.infer with a traced model and a TtaWrap dataset: same
    numbers as the reference order activation -> inverse -> numpy
    (utils/torch.py:24-29), but resident."""
    from mlcomp_b200.contrib.transform.tta import TtaWrap
    from mlcomp_b200.utils.torch import fuse, infer

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.conv = torch.nn.Conv2d(1, 2, 3, padding=1)

        def forward(self, x):
            return self.conv(x)

    torch.manual_seed(0)
    net = Net().eval()
    path = str(tmp_path / 'net.pth')
    torch.jit.save(torch.jit.trace(net, torch.zeros(1, 1, 8, 12)), path)
    data = torch.randn(7, 1, 8, 12)

    class DS(torch.utils.data.Dataset):
        def __init__(self, flip):
            self.flip = flip

        def __len__(self):
            return len(data)

        def __getitem__(self, i):
            x = data[i]
            return {'features': torch.flip(x, [2]) if self.flip else x}

    plain = infer(DS(False), path, batch_size=3, activation='sigmoid',
                  num_workers=0)
    flipped = infer(TtaWrap(DS(True), tfms=['hflip']), path, batch_size=3,
                    activation='sigmoid', num_workers=0)
    assert plain.tensor.is_cuda and flipped.tags == ['hflip']
    got = fuse([plain, flipped]).cpu()
    with torch.no_grad():
        a = torch.sigmoid(net(data))
        b = oracle.tta_inverse(torch.sigmoid(net(torch.flip(data, [3]))),
                               ['hflip'])
    want = (a + b) / 2
    assert torch.allclose(got, want, atol=2e-6)
    gen = infer(DS(False), path, batch_size=4, batch_mode=True, num_workers=0)
    counts = [b['count'] for b in gen]
    assert counts == [4, 3]


def test_add_host_with_run_length_ground_truth():
    """SegSearch.add_host(gt_runs=...): same scores as with dense planes."""
    from mlcomp_b200 import ops, scoring
    rng = np.random.default_rng(9)
    n, c, h, w = 11, 3, 40, 56
    gt = (rng.random((n, c, h, w)) < 0.2).astype(np.uint8)
    gt[2] = 0
    logits = torch.from_numpy((6 * (gt - 0.5) + rng.normal(0, 2, gt.shape))
                              .astype(np.float16))
    g = torch.from_numpy(gt)
    a = scoring.SegSearch(c, min_areas=(0, 50))
    a.add_host([logits], g, chunk=4)
    want = a.result()
    runs = tuple(t.cpu() for t in
                 ops.rle_encode([g.to(DEV)], 0.5, 0, activation=None)[:3])
    b = scoring.SegSearch(c, min_areas=(0, 50))
    b.add_host([logits], None, chunk=4, gt_runs=runs)
    got = b.result()
    assert np.array_equal(got['score'], want['score'])
    assert np.array_equal(got['micro_T'], want['micro_T'])
    assert got['n_img'] == n


def test_valid_segment_with_rle_targets(tmp_path, monkeypatch):
    """``targets_rle`` (csv of EncodedPixels, the competitions' at-rest form)
    gives the same search result as dense ``targets``."""
    import pandas as pd
    from mlcomp_b200.worker.executors import ValidSegment
    monkeypatch.chdir(tmp_path)
    rng = np.random.default_rng(17)
    n, c, h, w = 9, 2, 24, 40
    gt = (rng.random((n, c, h, w)) < 0.25).astype(np.uint8)
    gt[1] = 0
    logits = (6 * (gt - 0.5) + rng.normal(0, 2, gt.shape)).astype(np.float16)
    rows = [{'ImageId': f'{i}_{j + 1}',
             'EncodedPixels': oracle.rle_encode(gt[i, j])}
            for i in range(n) for j in range(c)]
    pd.DataFrame(rows).to_csv('train.csv', index=False)
    from mlcomp_b200.worker import store

    def run(ex):
        sess = store.MemorySession()
        dag = sess.add(store.Dag(name='d', img_size=0))
        task = sess.add(store.Task(name='v', dag=dag.id))
        ex.session = sess
        return ex(task=task, task_provider=sess, dag=dag)

    dense = ValidSegment(logits=[logits], targets=gt, part_size=4,
                         min_areas=[0, 20], plot_count=2,
                         imgs=np.zeros((n, h, w, 3), np.uint8), name='seg')
    want = run(dense)
    rle = ValidSegment(logits=[logits], targets_rle='train.csv', part_size=4,
                       min_areas=[0, 20], plot_count=2,
                       imgs=np.zeros((n, h, w, 3), np.uint8), name='seg')
    got = run(rle)
    assert got == want
    assert np.array_equal(rle.search.result()['score'],
                          dense.search.result()['score'])
