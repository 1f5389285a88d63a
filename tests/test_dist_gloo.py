"""N>1 host logic on CPU: world_size-2 gloo run of the shard split and of the
one exchange step (integer all-reduce + rank-ordered fp64 combine)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from mlcomp_b200.scoring import combine_ranks, shard_range
    n_total, c, b, a = 101, 3, 20, 2
    rng = np.random.default_rng(5)           # same stream on every rank
    per_img_dice = rng.random((n_total, c, b, a))
    per_img_micro = rng.integers(0, 1000, (n_total, c, 1 + 2 * b))
    lo, hi = shard_range(n_total, rank, world)
    dsum = torch.from_numpy(per_img_dice[lo:hi].sum(axis=0))
    micro = torch.from_numpy(per_img_micro[lo:hi].sum(axis=0))
    tot, m, n = combine_ranks(dsum, micro, hi - lo)
    # the packed form SegSearch exchanges: [dice | iou] fp64 bit patterns,
    # then the int64 counts and the image count - ONE collective
    from mlcomp_b200.scoring import combine_packed
    iou = torch.from_numpy((per_img_dice[lo:hi] ** 2).sum(axis=0))
    packed = torch.cat([dsum.reshape(-1).view(torch.int64),
                        iou.reshape(-1).view(torch.int64), micro.reshape(-1),
                        torch.tensor([hi - lo])])
    out = combine_packed(packed, 2 * dsum.numel())
    np.save(os.path.join(out_dir, f'packed_{rank}.npy'), out.numpy())
    np.save(os.path.join(out_dir, f'dice_{rank}.npy'), tot.numpy())
    np.save(os.path.join(out_dir, f'micro_{rank}.npy'), m.numpy())
    np.save(os.path.join(out_dir, f'n_{rank}.npy'), np.array([int(n), lo, hi]))
    dist.destroy_process_group()


def test_shard_and_combine_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world,
             join=True)
    rng = np.random.default_rng(5)
    per_img_dice = rng.random((101, 3, 20, 2))
    per_img_micro = rng.integers(0, 1000, (101, 3, 41))
    d = [np.load(tmp_path / f'dice_{r}.npy') for r in range(world)]
    m = [np.load(tmp_path / f'micro_{r}.npy') for r in range(world)]
    n = [np.load(tmp_path / f'n_{r}.npy') for r in range(world)]
    # shards tile [0, N) without overlap
    assert n[0][1] == 0 and n[0][2] == n[1][1] and n[1][2] == 101
    assert n[0][0] == n[1][0] == 101
    # integers exact, identical on both ranks
    assert np.array_equal(m[0], per_img_micro.sum(axis=0))
    assert np.array_equal(m[0], m[1])
    # fp64: bit-identical across ranks, equal to the rank-ordered sum
    assert np.array_equal(d[0], d[1])
    want = per_img_dice[:n[0][2]].sum(axis=0) + per_img_dice[n[0][2]:].sum(axis=0)
    assert np.array_equal(d[0], want)
    np.testing.assert_allclose(d[0], per_img_dice.sum(axis=0), rtol=1e-12)
    # packed buffer: same numbers in one exchange, identical on both ranks
    p = [np.load(tmp_path / f'packed_{r}.npy') for r in range(world)]
    assert np.array_equal(p[0], p[1])
    nd = d[0].size
    assert np.array_equal(p[0][:nd].view(np.float64), d[0].reshape(-1))
    sq = per_img_dice ** 2
    want_iou = sq[:n[0][2]].sum(axis=0) + sq[n[0][2]:].sum(axis=0)
    assert np.array_equal(p[0][nd:2 * nd].view(np.float64),
                          want_iou.reshape(-1))
    assert np.array_equal(p[0][2 * nd:-1], m[0].reshape(-1))
    assert p[0][-1] == 101


def test_shard_range_covers_everything():
    from mlcomp_b200.scoring import shard_range
    for n in (0, 1, 7, 6666, 30000):
        for w in (1, 2, 4, 8):
            edges = [shard_range(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


def test_combine_is_identity_without_process_group():
    from mlcomp_b200.scoring import combine_ranks
    d, m = torch.ones(2, 3, 1, dtype=torch.float64), torch.ones(2, 7).long()
    d2, m2, n2 = combine_ranks(d, m, 5)
    assert d2 is d and m2 is m and n2 == 5
