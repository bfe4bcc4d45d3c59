"""GPU: BASELINE.json configs 3, 4 and 5 at their named sizes through the C ABI, every reward compared with the CPU oracle
and the full tables (K, counts, bboxes, segment sizes) of 32 label maps and more; plus a seeded, time-boxed slice of the
randomised sweep in tools/fuzz_parity.py, so that the driver-run suite sees what the builder's own tools see.
Reference behaviour on these shapes: the per-image loop over any K (src/compute_rewards.cu:125-193) and masked,
overlapping bboxes (src/png_size_estimator.cu:88-117)."""
import os
import sys
from multiprocessing.pool import ThreadPool

import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5      # BASELINE.json north_star: sizes and rewards within 1e-5 relative; integers bit-exact
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_all(orc, images, labels, sizes, Kstride, S=1):
    """Oracle rewards + tables of every label map, the maps spread over the host cores (the C oracle releases the GIL)."""
    V = labels.shape[0]

    def one(v):
        r, det = orc.compute_rewards(images[v // S:v // S + 1], labels[v:v + 1], sizes[v // S:v // S + 1], details=True, Kstride=Kstride)
        return r[0], det["K"][0], det["counts"][0], det["bboxes"][0], det["seg_sizes"][0]
    with ThreadPool(min(32, os.cpu_count() or 1)) as pool:
        return pool.map(one, range(V))


def _compare(got_rewards, tables, want, n_tables):
    K, counts, bboxes, seg = tables
    worst = 0.0
    for v, (r, k, c, b, s) in enumerate(want):
        rel = abs(got_rewards[v] - r) / max(abs(r), 1e-300)
        worst = max(worst, rel)
        assert rel < RTOL, f"reward of label map {v}: {got_rewards[v]} vs {r}"
        assert K[v] == k, f"K of label map {v}"
        if v < n_tables:
            assert np.array_equal(counts[v, :k], c[:k]), f"counts of label map {v}"
            assert np.array_equal(bboxes[v, :k], b[:k]), f"bboxes of label map {v}"
            assert np.allclose(seg[v, :k], s[:k], rtol=RTOL, atol=0), f"segment sizes of label map {v}"
    return worst


def test_config3_full_size_256_x_512sq_voronoi(icr, orc, dev):
    """Config 3: batch 256 x 512x512 RGB, "SLIC-style" maps of 1024 segments.  64 distinct (image, map) pairs tiled four
    times: all 256 rewards are compared with the oracle (tiled copies bit for bit with their twin), the full tables of
    the 64 distinct maps too."""
    P, reps, H, W = 64, 4, 512, 512
    images, labels, sizes = synth.make_batch(3, B=P)
    want = _oracle_all(orc, images, labels, sizes, Kstride=1024)
    ws = icr.Workspace(dev, P * reps, H, W, 1024)
    ti = torch.from_numpy(images).to(dev).repeat(reps, 1, 1, 1)
    tl = torch.from_numpy(labels).to(dev).repeat(reps, 1, 1)
    ts = torch.from_numpy(sizes).to(dev).repeat(reps)
    got = icr.compute_rewards_batched(ti, tl, ts, workspace=ws).cpu().numpy()
    tables = [t.cpu().numpy() for t in ws.tables()]
    assert got.shape == (256,) and np.all(tables[0] == 1024)
    _compare(got, tables, want, n_tables=P)
    for r in range(1, reps):
        assert np.array_equal(got[:P], got[r * P:(r + 1) * P])
        for t in tables:
            assert np.array_equal(t[:P], t[r * P:(r + 1) * P], equal_nan=True)
    # the slices of the same call, sampled: Rule P bytes of overlapping bboxes at this size
    ws.inputs = (ti, tl, ts)
    buf, off = ws.slices()
    offh = off.cpu().numpy()
    rng = np.random.default_rng(3)
    for v in rng.choice(256, 6, replace=False):
        rgba = orc.chw3_f32_to_rgba(images[v % P])
        comp, uniq = orc.compact_labels(labels[v % P])
        for k in rng.choice(1024, 4, replace=False):
            x0, y0, x1, y1 = tables[2][v, k]
            w, h = x1 - x0 + 1, y1 - y0 + 1
            f = v * 1024 + k
            sl = buf[offh[f]:offh[f + 1]].cpu().numpy().reshape(h, w, 4)
            assert np.array_equal(sl, orc.slice_segment(rgba, comp, int(k), x0, y0, w, h))


def test_config4_full_size_256_x_224sq_8_maps_per_image(icr, orc, dev):
    """Config 4, the RL reward step: batch 256 x 224x224, 8 sampled partitions per image (BSP rectangles for even s,
    region growing for odd s), [B,S,H,W] labels with every image read once.  16 distinct images (128 distinct maps)
    tiled sixteen times; all 2048 rewards and the full tables of the 128 distinct maps against the oracle."""
    P, reps, S, H, W = 16, 16, 8, 224, 224
    images, labels, sizes = synth.make_multi_batch(B=P, S=S)
    flat = labels.reshape(P * S, H, W)
    Kmax = 512
    assert max(len(np.unique(m)) for m in flat) <= Kmax
    want = _oracle_all(orc, images, flat, sizes, Kstride=Kmax, S=S)
    ws = icr.Workspace(dev, P * reps * S, H, W, Kmax)
    ti = torch.from_numpy(images).to(dev).repeat(reps, 1, 1, 1)
    tl = torch.from_numpy(labels).to(dev).repeat(reps, 1, 1, 1)
    ts = torch.from_numpy(sizes).to(dev).repeat(reps)
    got = icr.compute_rewards_batched(ti, tl, ts, workspace=ws).cpu().numpy()
    assert got.shape == (256, 8)
    got = got.reshape(-1)
    tables = [t.cpu().numpy() for t in ws.tables()]
    _compare(got, tables, want, n_tables=P * S)
    n = P * S
    for r in range(1, reps):
        assert np.array_equal(got[:n], got[r * n:(r + 1) * n])
        assert np.array_equal(tables[0][:n], tables[0][r * n:(r + 1) * n])
    # the reference's own formulation of the same step: images.repeat_interleave(8, 0) through the S = 1 signature
    r1 = icr.compute_rewards_batched(ti[:P].repeat_interleave(S, 0), tl[:P].reshape(P * S, H, W), ts[:P].repeat_interleave(S)).cpu().numpy()
    assert np.array_equal(r1, got[:n])
    # Rule E over all 2048 rewards, then the standardised advantages (src/fcn/training.cpp:178-181)
    ema = icr.EMABaseline()
    value, adv = ema.update(torch.from_numpy(got).to(dev), advantages=True)
    assert abs(float(value) - got.mean()) <= 1e-12 * abs(got.mean())
    a = got - float(value)
    a = (a - a.mean()) / max(a.std(), 1e-6)
    assert np.allclose(adv.cpu().numpy(), a, rtol=1e-9, atol=1e-12)


def test_config5_shard_full_size_all_distinct_images(icr, orc, dev):
    """Config 5's per-GPU shard (512 x 224x224, what bench.py times): every one of the 64 distinct images of the pool
    against the oracle, rewards and full tables; the tiled copies bit for bit."""
    P, reps = 64, 8
    images, labels, sizes = synth.make_batch(5, B=P)
    want = _oracle_all(orc, images, labels, sizes, Kstride=256)
    ws = icr.Workspace(dev, P * reps, 224, 224, 256)
    ti = torch.from_numpy(images).to(dev).repeat(reps, 1, 1, 1)
    tl = torch.from_numpy(labels).to(dev).repeat(reps, 1, 1)
    ts = torch.from_numpy(sizes).to(dev).repeat(reps)
    got = icr.compute_rewards_batched(ti, tl, ts, workspace=ws).cpu().numpy()
    tables = [t.cpu().numpy() for t in ws.tables()]
    _compare(got, tables, want, n_tables=P)
    for r in range(1, reps):
        assert np.array_equal(got[:P], got[r * P:(r + 1) * P])
    # int64 labels, what compute_rewards_batched's own callers hold (src/compute_rewards.cu:97): same numbers
    got64 = icr.compute_rewards_batched(ti, tl.long(), ts, workspace=ws).cpu().numpy()
    assert np.array_equal(got64, got)


def test_background_segment_touching_all_borders(icr, orc, dev):
    """What real multicut output looks like: compact segments on a background that touches all four borders, so its bbox
    is the whole image and it takes the Rule M path (src/png_size_estimator.cu:397-463) while the rest do not."""
    B, H, W = 8, 224, 224
    images, labels, sizes = synth.make_batch(5, B=B, seed0=9100)
    for b in range(B):
        lab = labels[b]
        ids = np.unique(lab)
        rng = np.random.default_rng(b)
        bg = rng.choice(ids, size=len(ids) // 3, replace=False)           # a third of the rectangles melt into the background
        lab[np.isin(lab, bg)] = -5
        lab[0, :], lab[-1, :], lab[:, 0], lab[:, -1] = -5, -5, -5, -5
    want = _oracle_all(orc, images, labels, sizes, Kstride=256)
    ws = icr.Workspace(dev, B, H, W, 256)
    ti, tl, ts = (torch.from_numpy(a).to(dev) for a in (images, labels, sizes))
    got = icr.compute_rewards_batched(ti, tl, ts, workspace=ws).cpu().numpy()
    _compare(got, [t.cpu().numpy() for t in ws.tables()], want, n_tables=B)


def test_seeded_fuzz_slice(icr, orc, dev):
    """About 25 s of tools/fuzz_parity.py with a fixed seed: random shapes, seven label generators with scaled / offset /
    negative encodings in int32 or int64, five image generators, random parameters; tables exact, rewards at 1e-9,
    slices, slice plan and decoder round trip of every case."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import fuzz_parity
    cases, worst = fuzz_parity.run(10000, seed=20261105, budget_s=25.0, log=lambda *a: None)
    assert cases >= 20, f"only {cases} fuzz cases fitted the time box"
    assert worst < 1e-9


def test_big_segments_shared_among_ctas_give_the_same_sizes(icr, orc, dev):
    """A batch with fewer big segments than half the grid shares each of them out among several CTAs (bands of rows, merged
    through integer atomics); a large batch gives every big segment one CTA.  Same maps, both ways: sizes and rewards bit
    for bit, and both against the oracle.  One whole-image segment per map plus maps with a background around rectangles."""
    H, W = 224, 224
    images, labels, sizes = synth.make_batch(5, B=8, seed0=7300)
    labels = labels.copy()
    labels[:4] = 9                                                        # whole-image segments (Rule M, one-pixel chunks)
    for b in range(4, 8):                                                 # rectangles on a border-touching background
        labels[b][0, :], labels[b][-1, :], labels[b][:, 0], labels[b][:, -1] = -5, -5, -5, -5
    want = _oracle_all(orc, images, labels, sizes, Kstride=256)
    small = icr.Workspace(dev, 8, H, W, 256)
    ti, tl, ts = (torch.from_numpy(a).to(dev) for a in (images, labels, sizes))
    got_small = icr.compute_rewards_batched(ti, tl, ts, workspace=small).cpu().numpy()
    tab_small = [t.cpu().numpy() for t in small.tables()]
    _compare(got_small, tab_small, want, n_tables=8)
    reps = 40                                                             # 320 maps: more than half the grid, one CTA per big segment
    large = icr.Workspace(dev, 8 * reps, H, W, 256)
    got_large = icr.compute_rewards_batched(ti.repeat(reps, 1, 1, 1), tl.repeat(reps, 1, 1), ts.repeat(reps), workspace=large).cpu().numpy()
    tab_large = [t.cpu().numpy() for t in large.tables()]
    assert np.array_equal(got_large[:8], got_small) and np.array_equal(got_large[-8:], got_small)
    for v in range(8):                                                    # segment sizes of the rows in use (rows past K are never written)
        k = tab_small[0][v]
        assert tab_large[0][v] == k and np.array_equal(tab_large[3][v, :k], tab_small[3][v, :k])
    # and again on the small workspace: the merge slots were left clear
    again = icr.compute_rewards_batched(ti, tl, ts, workspace=small).cpu().numpy()
    assert np.array_equal(again, got_small)
