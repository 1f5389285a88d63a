"""GPU parity of seg_hist_guarded (csrc/seg_hist_guarded.cu): the multi-source
fp16 sigmoid histogram that decides most pixels from a bounded-error estimate
and the rest exactly.  Its results must equal, bit for bit, the binning of the
exactly fused probabilities (mlc_seg_prob, itself bit-equal to torch's CUDA
sigmoid + the oracle's fp32 mean): histograms AND per-pixel bin planes.
Reference lines behind the arithmetic: utils/torch.py:15,24-29 (sigmoid on the
device, then TtaWrap.inverse), contrib/transform/tta.py:20-31,
contrib/metrics/dice.py:6-11 (`>=`, the sums)."""
import numpy as np
import pytest
import torch

import oracle
from test_gpu_parity import DEV, forward_tta, synth_seg

pytestmark = pytest.mark.gpu

FLIPS = [[], ['hflip'], ['vflip'], ['hflip', 'vflip']]


@pytest.fixture(scope='module')
def ops():
    from mlcomp_b200 import ops as o
    return o


def device_bins(ops, srcs, tags, thr):
    """cnt(p) of the exactly fused probabilities, on the device:
    #{k : p >= thr[k]} = searchsorted(thr, p, right=True); NaN -> 0."""
    p = ops.seg_prob(srcs, xforms=tags)
    t = torch.from_numpy(np.asarray(thr, np.float32)).to(p.device)
    c = torch.searchsorted(t, p.reshape(-1), right=True).reshape(p.shape)
    c[torch.isnan(p)] = 0
    return c


def check(ops, srcs, tags, gt, thr):
    hist, bins = ops.seg_hist(srcs, gt, thr, xforms=tags, bins=True)
    hist2 = ops.seg_hist(srcs, gt, thr, xforms=tags)
    assert torch.equal(hist, hist2)
    want = device_bins(ops, srcs, tags, thr)
    assert torch.equal(bins.long(), want), \
        int((bins.long() != want).sum())
    nb = len(thr) + 1
    n, c = gt.shape[:2]
    flat = want.reshape(n * c, -1) + nb * (gt.reshape(n * c, -1) != 0).long()
    ref = torch.zeros((n * c, 2 * nb), dtype=torch.long, device=gt.device)
    ref.scatter_add_(1, flat, torch.ones_like(flat))
    assert torch.equal(hist.reshape(n * c, 2 * nb).long(), ref)
    return hist


@pytest.mark.parametrize('n_src', [2, 3, 4, 5, 8, 9])
@pytest.mark.parametrize('flips', [False, True])
def test_guarded_matches_exact_binning(ops, n_src, flips):
    srcs, gt = synth_seg(3, 2, 64, 96, seed=300 + n_src, n_src=n_src)
    tags = [FLIPS[k % 4] if flips else [] for k in range(n_src)]
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, tags)]
    thr = oracle.threshold_grid(100)
    hist = check(ops, seen, tags, gt.to(DEV), thr)
    # and against the oracle's own binning of those probabilities on the CPU
    p = ops.seg_prob(seen, xforms=tags).cpu().numpy()
    for i in range(gt.shape[0]):
        for j in range(gt.shape[1]):
            assert np.array_equal(
                hist[i, j].cpu().numpy(),
                oracle.bin_counts(gt[i, j].numpy(), p[i, j], thr))


def test_guarded_every_fp16_code(ops):
    """All 65 536 codes (NaN, +-inf, +-65504, denormals, +-0) as S equal
    sources and against shifted copies: overflow of the denominator product
    and NaN items must take the exact path."""
    codes = torch.arange(65536, dtype=torch.int32).to(torch.int16)
    x = codes.view(torch.float16).reshape(1, 1, 256, 256).to(DEV)
    gt = (torch.arange(65536).reshape(1, 1, 256, 256) % 3 == 0) \
        .to(torch.uint8).to(DEV)
    thr = oracle.threshold_grid(100)
    for n_src in (2, 4, 5):
        check(ops, [x] * n_src, [[]] * n_src, gt, thr)
    rolled = [x, x.flatten().roll(4099).reshape(x.shape),
              x.flatten().roll(17).reshape(x.shape),
              x.flatten().roll(33333).reshape(x.shape)]
    rolled = [r.contiguous() for r in rolled]
    check(ops, rolled, [[]] * 4, gt, thr)
    check(ops, rolled, FLIPS, gt, thr)


def test_guarded_means_on_and_next_to_thresholds(ops):
    """Sources whose exact mean lands on a threshold or 1 ulp beside it: the
    estimate cannot decide those pixels, the exact path must."""
    thr = oracle.threshold_grid(100)
    halves = torch.arange(-0x4C00, 0x4C00, dtype=torch.int32)
    vals = torch.where(halves < 0, (-halves) | 0x8000, halves).to(torch.int16) \
        .view(torch.float16)                      # every |x| < 16
    sig = torch.sigmoid(vals.float().to(DEV))     # the device's exact sigmoid
    t = torch.from_numpy(thr).to(DEV)
    # for every threshold: pairs (x, x') with (s(x) + s(x')) / 2 nearest to it
    rows = []
    for k in range(1, 100):
        target = 2 * t[k] - sig[::37]             # wanted s(x')
        j = torch.searchsorted(sig, target).clamp(1, len(sig) - 2)
        for d in (-1, 0, 1):
            rows.append(torch.stack([vals[::37].to(DEV), vals.to(DEV)[j + d]]))
    pairs = torch.cat(rows, dim=1)                # [2, n]
    n = pairs.shape[1] // 8 * 8
    a = pairs[0, :n].reshape(1, 1, 1, n).contiguous()
    b = pairs[1, :n].reshape(1, 1, 1, n).contiguous()
    gt = torch.zeros((1, 1, 1, n), dtype=torch.uint8, device=DEV)
    check(ops, [a, b], [[], []], gt, thr)
    check(ops, [a, b, b, a], [[], []] * 2, gt, thr)


@pytest.mark.parametrize('thr', [
    np.float32([0.5]),                                  # one threshold
    np.float32([0.1, 0.11, 0.5, 0.9]),                  # irregular
    np.float32(np.arange(50) / 100 + 0.25),             # does not span [0, 1]
    np.float32(np.arange(200) / 200),                   # more than 120
    np.float32(np.arange(10) / 10),                     # coarse uniform
], ids=['one', 'irregular', 'partial', 'fine', 'coarse'])
def test_guarded_or_fallback_any_table(ops, thr):
    """Grids the guarded kernel declines stay on the exact kernels; either
    way the result is the exact binning."""
    srcs, gt = synth_seg(2, 2, 32, 64, seed=41, n_src=4)
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, FLIPS)]
    check(ops, seen, FLIPS, gt.to(DEV), thr)


def test_guarded_severstal_tta_full_planes(ops):
    """4-flip TTA on 256x1600 planes (26 M pixels): every pixel's bin equals
    the binning of the exact probabilities; the dice table equals the one
    computed from the exact histograms."""
    g = torch.Generator(device=DEV).manual_seed(5)
    n, c, h, w = 16, 4, 256, 1600
    base = torch.randn((n, c, h, w), generator=g, device=DEV) * 1.5 - 4
    base[:, :, 64:128, 200:900] += 8
    gt = torch.zeros((n, c, h, w), dtype=torch.uint8, device=DEV)
    gt[:, :, 64:128, 200:900] = 1
    gt[::2] = 0
    srcs = []
    for dims in ((), (3,), (2,), (2, 3)):
        v = (base + 0.5 * torch.randn(base.shape, generator=g, device=DEV)) \
            .half()
        srcs.append(v.flip(*dims).contiguous() if dims else v)
    thr = oracle.threshold_grid(100)
    check(ops, srcs, FLIPS, gt, thr)
    # a sharper set: confident logits N(+-6, 1)
    sharp = [(s.float() * 0.66 + (s.float() > 0) * 4 - 2).half() for s in srcs]
    check(ops, sharp, FLIPS, gt, thr)


def test_bins_into_a_callers_buffer(ops):
    """seg_hist(bins=<tensor>) / SegSearch.add(keep_bins=<tensor>) write the
    bin planes into the caller's buffer (no allocation per chunk) and give
    what bins=True gives; a buffer of another shape or dtype is refused."""
    from mlcomp_b200 import scoring
    srcs, gt = synth_seg(3, 2, 16, 64, seed=11, n_src=4)
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, FLIPS)]
    thr = oracle.threshold_grid(100)
    hist, bins = ops.seg_hist(seen, gt.to(DEV), thr, xforms=FLIPS, bins=True)
    buf = torch.full((3, 2, 16, 64), 255, dtype=torch.uint8, device=DEV)
    hist2, bins2 = ops.seg_hist(seen, gt.to(DEV), thr, xforms=FLIPS, bins=buf)
    assert bins2.data_ptr() == buf.data_ptr()
    assert torch.equal(bins2, bins) and torch.equal(hist2, hist)
    s = scoring.SegSearch(2, thr, [0, 8])
    buf.fill_(255)
    h3, b3 = s.add(seen, gt.to(DEV), xforms=FLIPS, keep_bins=buf)
    assert b3.data_ptr() == buf.data_ptr() and torch.equal(b3, bins)
    assert torch.equal(h3, hist)
    assert isinstance(s.add(seen, gt.to(DEV), xforms=FLIPS), torch.Tensor)
    with pytest.raises(ValueError):
        ops.seg_hist(seen, gt.to(DEV), thr, xforms=FLIPS, bins=buf[:2])
    with pytest.raises(TypeError):
        ops.seg_hist(seen, gt.to(DEV), thr, xforms=FLIPS, bins=buf.int())


def test_guarded_split_planes_small_job(ops):
    """Few large planes: the host splits planes into units (atomic flush)."""
    g = torch.Generator(device=DEV).manual_seed(9)
    x = [(torch.randn((1, 1, 1024, 2048), generator=g, device=DEV) * 3).half()
         for _ in range(4)]
    gt = (torch.rand((1, 1, 1024, 2048), generator=g, device=DEV) < 0.3) \
        .to(torch.uint8)
    check(ops, x, [[]] * 4, gt, oracle.threshold_grid(100))
    check(ops, x, FLIPS, gt, oracle.threshold_grid(100))


# ------------------------------------------------- pipeline: known-area RLE
def test_rle_known_area_skips_only_empty_masks(ops):
    """mlc_rle_encode_known with areas from the histograms == mlc_rle_encode,
    for the bin planes and for the sources; strings equal the oracle's."""
    srcs, gt = synth_seg(6, 3, 64, 96, seed=71, n_src=4, empty=0.6)
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, FLIPS)]
    thr = oracle.threshold_grid(100)
    hist, bins = ops.seg_hist(seen, gt.to(DEV), thr, xforms=FLIPS, bins=True)
    p = ops.seg_prob(seen, xforms=FLIPS).cpu().numpy()
    for k_c, a_c in (([50, 70, 98], [0, 30, 0]), ([10, 99, 0], [5000, 1, 3])):
        t_c = [float(thr[k]) for k in k_c]
        area = ops.plane_area(hist, k_c)
        want_area = torch.from_numpy(
            np.stack([[(p[i, j] >= thr[k_c[j]]).sum() for j in range(3)]
                      for i in range(6)])).int()
        assert torch.equal(area.cpu(), want_area)
        a = ops.rle_from_bins(bins, k_c, a_c, hist=hist)
        b = ops.rle_from_bins(bins, k_c, a_c)
        c = ops.rle_encode(seen, t_c, a_c, xforms=FLIPS, known_area=area)
        d = ops.rle_encode(seen, t_c, a_c, xforms=FLIPS)
        sa, sb, sc, sd = (ops.rle_strings(*x[:3]) for x in (a, b, c, d))
        assert sa == sb == sc == sd
        assert torch.equal(a[3], d[3]) and torch.equal(c[3], d[3])
        k = 0
        for i in range(6):
            for j in range(3):
                assert sa[k] == oracle.rle_submission(p[i, j], t_c[j], a_c[j])
                k += 1
        assert any(s == '' for s in sa) and any(s != '' for s in sa)


def test_add_host_twice_without_sync(ops):
    """Two add_host calls with different data and no synchronisation in
    between must not overwrite staging buffers still being read."""
    from mlcomp_b200 import scoring
    s1, g1 = synth_seg(9, 2, 64, 96, seed=5, n_src=2)
    s2, g2 = synth_seg(9, 2, 64, 96, seed=6, n_src=2)
    tags = [[], ['hflip']]
    a = scoring.SegSearch(2, min_areas=[0, 50])
    a.reset()
    a.add([s.to(DEV) for s in s1], g1.to(DEV), tags)
    a.add([s.to(DEV) for s in s2], g2.to(DEV), tags)
    b = scoring.SegSearch(2, min_areas=[0, 50])
    b.reset()
    # same host buffers reused for both sets: the second set is written into
    # them only after the first call's copies were issued
    hs = [s.clone().pin_memory() for s in s1]
    hg = g1.clone().pin_memory()
    b.add_host(hs, hg, tags, chunk=2)
    torch.cuda.current_stream().synchronize()    # host buffers reusable
    for d, s in zip(hs, s2):
        d.copy_(s)
    hg.copy_(g2)
    b.add_host(hs, hg, tags, chunk=2)
    b.add_host(hs, hg, tags, chunk=2)            # and once more, back to back
    ra, rb = a.result(), b.result()
    a2 = scoring.SegSearch(2, min_areas=[0, 50])
    a2.reset()
    a2.add([s.to(DEV) for s in s1], g1.to(DEV), tags)
    a2.add([s.to(DEV) for s in s2], g2.to(DEV), tags)
    a2.add([s.to(DEV) for s in s2], g2.to(DEV), tags)
    ra = a2.result()
    assert np.array_equal(ra['micro_P'], rb['micro_P'])
    assert np.array_equal(ra['micro_I'], rb['micro_I'])
    np.testing.assert_allclose(ra['score'], rb['score'], rtol=1e-12)


def test_valid_segment_prepare_submit_from_host_arrays(ops, tmp_path):
    """The executor path the bench's e2e takes: host arrays (pinned tensors,
    and the same data as memory-mapped .npy files) staged part by part,
    search, then the run-length csv at the chosen point from the kept bin
    planes - against the device-resident pipeline and the oracle."""
    from mlcomp_b200 import scoring
    from mlcomp_b200.worker import store
    from mlcomp_b200.worker.executors import Executor
    srcs, gt = synth_seg(10, 2, 64, 96, seed=91, n_src=4, empty=0.4)
    seen = [forward_tta(s, t).contiguous() for s, t in zip(srcs, FLIPS)]
    paths = []
    for k, s in enumerate(seen):
        paths.append(str(tmp_path / f'v{k}.npy'))
        np.save(paths[-1], s.numpy())
    np.save(str(tmp_path / 'gt.npy'), gt.numpy())
    areas = [0, 40, 400]
    ref = scoring.SegSearch(2, min_areas=areas)
    ref.reset()
    hist, bins = ref.add([s.to(DEV) for s in seen], gt.to(DEV), FLIPS,
                         keep_bins=True)
    want = ref.result()
    k_c = [b['threshold_index'] for b in want['best']]
    a_c = [b['min_area'] for b in want['best']]
    want_strings = ops.rle_strings(*ops.rle_from_bins(bins, k_c, a_c)[:3])
    for logits, targets in (
            ([s.pin_memory() for s in seen], gt.pin_memory()),
            (paths, str(tmp_path / 'gt.npy'))):
        cfg = {'executors': {'valid': {
            'type': 'valid_segment', 'logits': logits, 'tta': FLIPS,
            'targets': targets, 'min_areas': areas, 'part_size': 3,
            'prepare_submit': True}}}
        ex = Executor.from_config(executor='valid', config=cfg,
                                  additional_info={})
        out = ex(task=store.Task(name='t'), task_provider=None,
                 dag=store.Dag(name='d'))
        assert abs(out['res'] - want['dice']) < 1e-12
        assert [b['threshold_index'] for b in ex.best] == k_c
        assert [r['EncodedPixels'] for r in ex.rows] == want_strings
        assert ex.rows[3]['ImageId'] == '1_2'
        assert ex.d2h_bytes > 0


# ------------------------------------------- rows of any width / alignment
@pytest.mark.parametrize('shape', [(3, 2, 35, 53), (2, 4, 350, 525),
                                   (2, 1, 9, 16), (1, 2, 40, 23),
                                   (5, 3, 7, 100)])
@pytest.mark.parametrize('n_src,flips', [(4, True), (2, True), (3, False),
                                         (4, False)])
def test_guarded_unaligned_rows(ops, shape, n_src, flips):
    """350 x 525-like planes (rows that are no multiple of 16 bytes, planes
    that start at any element offset; csrc/seg_hist_guarded_rows.cu also for
    rows of a few items): the exact binning."""
    n, c, h, w = shape
    srcs, gt = synth_seg(n, c, h, w, seed=h + w + n_src, n_src=n_src)
    tags = [FLIPS[k % 4] if flips else [] for k in range(n_src)]
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, tags)]
    check(ops, seen, tags, gt.to(DEV), oracle.threshold_grid(100))


@pytest.mark.parametrize('shape', [(3, 3, 9, 263), (5, 4, 4, 250),
                                   (2, 9, 3, 257), (1, 1, 6, 249),
                                   (1, 2, 3, 601), (1, 1, 2, 1001),
                                   (2, 2, 40, 525), (1, 33, 2, 255)])
@pytest.mark.parametrize('n_src,flips', [(4, True), (2, True), (3, False)])
def test_guarded_tma_rows(ops, shape, n_src, flips):
    """Wide rows of any width (csrc/seg_hist_guarded_rows.cu: bands of rows
    staged by bulk copies): partial last items, one to many rows per band,
    more bands than ring slots, planes at every 16-byte phase."""
    n, c, h, w = shape
    srcs, gt = synth_seg(n, c, h, w, seed=h + w + n_src, n_src=n_src)
    tags = [FLIPS[k % 4] if flips else [] for k in range(n_src)]
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, tags)]
    check(ops, seen, tags, gt.to(DEV), oracle.threshold_grid(100))


@pytest.mark.parametrize('seed', range(24))
def test_guarded_rows_random_geometry(ops, seed):
    """Random widths (9 .. 700, mostly no multiple of 8), heights, plane
    counts, source counts and flip sets through the band-staging kernel:
    histograms and bin planes equal the exact binning."""
    rng = np.random.default_rng(1000 + seed)
    w = int(rng.integers(9, 700))
    if seed % 6 == 0:
        w = w // 8 * 8 + 8          # rows of whole items at odd plane starts
    h = int(rng.integers(1, 24))
    n, c = int(rng.integers(1, 4)), int(rng.integers(1, 6))
    n_src = int(rng.integers(2, 5))
    tags = [FLIPS[int(k)] for k in rng.integers(0, 4, n_src)]
    srcs, gt = synth_seg(n, c, h, w, seed=seed, n_src=n_src)
    seen = [forward_tta(s_, t) for s_, t in zip(srcs, tags)]
    # planes that start at any 16-byte phase: views into a larger buffer
    dev = []
    for s_ in seen:
        pad = int(rng.integers(0, 4)) * 8
        buf = torch.empty(s_.numel() + pad, dtype=torch.float16, device=DEV)
        v = buf[pad:].view(s_.shape)
        v.copy_(s_)
        dev.append(v)
    check(ops, dev, tags, gt.to(DEV), oracle.threshold_grid(100))


def test_guarded_tma_rows_special_values_and_tensor_ends(ops):
    """NaN / inf / huge values next to row ends, and tensors that end exactly
    where their allocation ends (a copy that runs over the end of the tensor
    must be zero-filled by the copy engine, not read)."""
    g = torch.Generator().manual_seed(5)
    n, c, h, w = 1, 3, 11, 253
    base = torch.randn((n, c, h, w), generator=g) * 4
    base[0, 0, 3, 252] = float('nan')
    base[0, 0, 4, 0] = float('inf')
    base[0, 1, 0, 0] = float('-inf')
    base[0, 2, 10, 252] = 65504.0
    base[0, 2, 10, 248] = float('nan')
    base[0, 1, 10, 240:248] = -30.0
    srcs = [(base + 0.25 * k).half() for k in range(4)]
    gt = (torch.rand((n, c, h, w), generator=g) < 0.3).to(torch.uint8)
    seen = []
    for s, t in zip(srcs, FLIPS):
        f = forward_tta(s, t)
        # the tensor's last byte is the allocation's last byte (512-byte
        # granules of the caching allocator: pick the offset accordingly)
        numel = f.numel()
        granule = 256
        total = (numel + granule - 1) // granule * granule
        buf = torch.empty(total, dtype=torch.float16, device=DEV)
        off = (total - numel) // 8 * 8
        view = buf[off:off + numel].view(f.shape)
        view.copy_(f)
        seen.append(view)
    check(ops, seen, FLIPS, gt.to(DEV), oracle.threshold_grid(100))


def test_guarded_unaligned_special_values(ops):
    """NaN / inf / huge values in windows that straddle vectors and rows."""
    g = torch.Generator().manual_seed(3)
    n, c, h, w = 1, 2, 21, 37
    base = torch.randn((n, c, h, w), generator=g) * 4
    base[0, 0, 3, 5] = float('nan')
    base[0, 0, 4, 36] = float('inf')
    base[0, 1, 0, 0] = float('-inf')
    base[0, 1, 20, 36] = 65504.0
    base[0, 1, 10, 8:16] = -30.0
    srcs = [(base + 0.25 * k).half() for k in range(4)]
    seen = [forward_tta(s, t).to(DEV) for s, t in zip(srcs, FLIPS)]
    gt = (torch.rand((n, c, h, w), generator=g) < 0.3).to(torch.uint8)
    check(ops, seen, FLIPS, gt.to(DEV), oracle.threshold_grid(100))
