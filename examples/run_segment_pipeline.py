#!/usr/bin/env python
"""Runs examples/segment_pipeline.yml on synthetic data, without MLComp's
server: the two executors are built with ``Executor.from_config`` and called
the way ``worker/tasks.py`` calls them (the valid task's result reaches the
infer task through ``task.result``, tasks.py:233-235 / executor.py:222-227).

Needs a CUDA device:  python examples/run_segment_pipeline.py [workdir]
"""
import os
import sys
import tempfile

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mlcomp_b200.contrib.transform.rle import masks2rle  # noqa: E402
from mlcomp_b200.worker import store  # noqa: E402
from mlcomp_b200.worker.executors import Executor  # noqa: E402


def synth(n, c, h, w, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    gt = np.zeros((n, c, h, w), np.uint8)
    for i in range(n):
        for j in range(c):
            if rng.random() < 0.5:
                continue
            cy, cx = rng.integers(0, h), rng.integers(0, w)
            ry, rx = rng.integers(4, h // 3), rng.integers(4, w // 3)
            gt[i, j] = ((yy - cy) / ry) ** 2 + ((xx - cx) / rx) ** 2 <= 1
    base = 8 * (gt.astype(np.float32) - 0.5) + rng.normal(0, 1.5, gt.shape)
    v0 = base.astype(np.float16)
    # the second variant is what a model fed the h-flipped image returns
    v1 = (base + rng.normal(0, 0.5, gt.shape))[..., ::-1].astype(np.float16)
    return gt, v0, np.ascontiguousarray(v1)


def main():
    work = sys.argv[1] if len(sys.argv) > 1 else tempfile.mkdtemp()
    os.chdir(work)
    n, c, h, w = 128, 4, 64, 160
    gt, v0, v1 = synth(n, c, h, w, 1)
    np.save('v0.npy', v0)
    np.save('v1.npy', v1)
    import pandas as pd
    import torch
    # the ground truth as the competitions ship it: one RLE string per plane
    strings = masks2rle(torch.from_numpy(gt).cuda())
    pd.DataFrame([{'ImageId': f'{k // c}_{k % c + 1}', 'EncodedPixels': s}
                  for k, s in enumerate(strings)]
                 ).to_csv('train.csv', index=False)
    _, t0, t1 = synth(n, c, h, w, 2)
    np.save('t0.npy', t0)
    np.save('t1.npy', t1)

    config = yaml.safe_load(open(os.path.join(ROOT, 'examples',
                                              'segment_pipeline.yml')))
    sess = store.MemorySession()
    dag = sess.add(store.Dag(name='demo', img_size=0))
    t_valid = sess.add(store.Task(name='valid', dag=dag.id))
    t_infer = sess.add(store.Task(name='infer', dag=dag.id))
    t_valid._child_of = t_infer.id          # infer depends on valid

    valid = Executor.from_config(executor='valid', config=config,
                                 additional_info=None, session=sess)
    valid.imgs = np.zeros((n, h, w, 3), np.uint8)
    res = valid(task=t_valid, task_provider=sess, dag=dag)
    t_valid.result = yaml.safe_dump(res)
    print('valid:', {k: res[k] for k in ('res', 'thresholds', 'min_areas')})

    infer = Executor.from_config(executor='infer', config=config,
                                 additional_info=None, session=sess)
    infer(task=t_infer, task_provider=sess, dag=dag)
    sub = pd.read_csv('data/submissions/unet_test.csv', keep_default_na=False)
    print('infer:', len(sub), 'rows in',
          os.path.join(work, 'data/submissions/unet_test.csv'))
    print(sub.head(3).to_string())


if __name__ == '__main__':
    main()
