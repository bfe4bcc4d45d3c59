#!/usr/bin/env python
"""Randomised parity sweep: the CUDA path (through the Python mirror of the C ABI) against the CPU oracle on random shapes,
partitions, label encodings and parameters.  `python tools/fuzz_parity.py [iterations] [seed]` on a GPU box runs for
minutes and exits non-zero on the first mismatch, printing the case's recipe; tests/test_full_size_configs.py runs a
seeded, time-boxed slice of the same sweep (`run`) inside the GPU suite."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def random_labels(rng, H, W):
    kind = rng.integers(0, 7)
    yy, xx = np.mgrid[0:H, 0:W]
    if kind == 0:
        lab = np.zeros((H, W), np.int64)
    elif kind == 1:                                            # noise: many tiny segments, long label runs are rare
        lab = rng.integers(0, max(2, (H * W) // int(rng.integers(2, 40))), (H, W))
    elif kind == 2:
        lab = (xx // int(rng.integers(1, 9))) + 1000 * (yy // int(rng.integers(1, 9)))
    elif kind == 3:
        from image_compression_cpp_b200 import synth
        lab = synth.bfs_partition(H, W, int(rng.integers(1, 30)), seed=int(rng.integers(1 << 30))) if H * W >= 64 else np.zeros((H, W), np.int64)
    elif kind == 4:
        from image_compression_cpp_b200 import synth
        lab = synth.voronoi_partition(H, W, cell=int(rng.integers(3, 24)), jitter=int(rng.integers(0, 8)), seed=int(rng.integers(1 << 30)))
    elif kind == 5:                                            # diagonal bands: bboxes overlap heavily
        lab = (xx + yy) // int(rng.integers(1, 12))
    else:                                                      # a few random rectangles painted over a background
        lab = np.zeros((H, W), np.int64)
        for i in range(int(rng.integers(1, 12))):
            y0, x0 = int(rng.integers(0, H)), int(rng.integers(0, W))
            lab[y0:y0 + int(rng.integers(1, H + 1)), x0:x0 + int(rng.integers(1, W + 1))] = i + 1
    scale, off = int(rng.choice([1, 1, 3, 7, 1 << 8])), int(rng.choice([0, 0, 5, -1000, 1 << 20]))
    return lab.astype(np.int64) * scale + off


def random_image(rng, H, W):
    kind = rng.integers(0, 5)
    if kind == 0:
        img = rng.random((3, H, W), dtype=np.float32) * 255.0
    elif kind == 1:
        img = rng.integers(0, 256, (3, H, W)).astype(np.float32)
    elif kind == 2:                                            # smooth ramps with halves (round-half-even cases)
        yy, xx = np.mgrid[0:H, 0:W]
        img = np.stack([xx * 1.5, yy * 2.5, (xx + yy) * 0.5]).astype(np.float32)
    elif kind == 3:
        img = rng.random((3, H, W), dtype=np.float32)          # the [0,1] quirk: everything collapses to 0/1
    else:
        img = (rng.standard_normal((3, H, W)) * 200.0 + 100.0).astype(np.float32)   # values beyond 0..255: clamped
    return np.ascontiguousarray(img, dtype=np.float32)


class Mismatch(AssertionError):
    pass


def run(iters, seed, budget_s=None, log=print):
    """Runs up to `iters` random cases (or until `budget_s` seconds have passed); returns (cases run, worst relative
    reward error).  Raises Mismatch with the case's recipe on the first disagreement with the oracle."""
    import torch
    from image_compression_cpp_b200 import reward as icr
    from oracle import oracle as orc
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda:0")
    t0, worst = time.time(), 0.0
    done = 0
    for it in range(iters):
        if budget_s is not None and time.time() - t0 > budget_s:
            break
        done += 1
        H = int(rng.choice([1, 2, 3, 5, 8, 16, 24, 31, 48, 64, 100, 128, 160]))
        W = int(rng.choice([1, 3, 4, 7, 8, 12, 16, 20, 28, 32, 36, 40, 56, 64, 100, 128, 200, 256, 300]))
        B, S = int(rng.integers(1, 4)), int(rng.choice([1, 1, 2]))
        images = np.stack([random_image(rng, H, W) for _ in range(B)])
        labels = np.stack([random_labels(rng, H, W) for _ in range(B * S)])
        dtype = np.int32 if (rng.random() < 0.6 and np.abs(labels).max() < 2 ** 31) else np.int64
        labels = labels.astype(dtype)
        sizes = rng.random(B) * 1e5 + 100.0
        kw = dict(min_pixels_per_segment=int(rng.choice([1, 1, 5, 50])), L_min=int(rng.choice([1, 2, 4, 4, 8])),
                  adaptive_filter=bool(rng.random() < 0.8), lambda_=float(rng.choice([0.5, 0.0, 2.0])))
        recipe = dict(it=it, seed=seed, H=H, W=W, B=B, S=S, dtype=dtype.__name__, **kw)
        prm = orc.params(kw["min_pixels_per_segment"], kw["L_min"], 0.012167, 18.0, 0.1, 9.308622, kw["adaptive_filter"], kw["lambda_"])
        want, det = orc.compute_rewards(np.repeat(images, S, axis=0), labels, np.repeat(sizes, S), prm, details=True, Kstride=H * W)
        ws = icr.Workspace(dev, B * S, H, W, None)
        ti, ts = torch.from_numpy(images).to(dev), torch.from_numpy(sizes).to(dev)
        tl = torch.from_numpy(labels.reshape((B, S, H, W) if S > 1 else (B, H, W))).to(dev)
        cap = icr.lib().icr_label_capacity(H, W)
        if any(int(m.max()) - int(m.min()) >= cap for m in labels):
            # beyond the C ABI's contract (spread below the bitmap capacity): it must say so; the host layers rank first
            try:
                icr.compute_rewards_batched(ti, tl, ts, workspace=ws, **kw)
                raise Mismatch(f"MISSING ERROR {recipe}")
            except RuntimeError as e:
                assert "label spread" in str(e), e
            ranked = np.stack([np.unique(m, return_inverse=True)[1].reshape(H, W) for m in labels]).astype(dtype)
            tl = torch.from_numpy(ranked.reshape((B, S, H, W) if S > 1 else (B, H, W))).to(dev)
        got = icr.compute_rewards_batched(ti, tl, ts, workspace=ws, **kw).cpu().numpy().reshape(-1)
        K, counts, bboxes, seg = [t.cpu().numpy() for t in ws.tables()]
        ok = np.array_equal(K[:B * S], det["K"])
        for v in range(B * S):
            k = det["K"][v]
            ok = ok and np.array_equal(counts[v, :k], det["counts"][v, :k]) and np.array_equal(bboxes[v, :k], det["bboxes"][v, :k])
            ok = ok and np.allclose(seg[v, :k], det["seg_sizes"][v, :k], rtol=1e-9, atol=0)
        rel = float(np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)))
        worst = max(worst, rel)
        if not ok or rel > 1e-9:
            raise Mismatch(f"MISMATCH {recipe} rel {rel} tables ok {ok}")
        # the slices of the same call: a few of them against the oracle's Rule P, and the decoder round trip of map 0
        ws.inputs = (ti, tl, ts)
        swap = bool(rng.integers(0, 2))
        buf, off = ws.slices(order=int(swap))
        keys = ws.keys().cpu().numpy()
        bufh, offh = buf.cpu().numpy(), off.cpu().numpy()
        maps = tl.cpu().numpy().reshape(B * S, H, W)
        f0 = 0
        for v in range(B * S):
            rgba = orc.chw3_f32_to_rgba(images[v // S])
            comp, uniq = orc.compact_labels(maps[v])
            for k in rng.choice(det["K"][v], size=min(3, det["K"][v]), replace=False):
                x0, y0, x1, y1 = det["bboxes"][v, k]
                w, h = x1 - x0 + 1, y1 - y0 + 1
                sl = bufh[offh[f0 + k]:offh[f0 + k + 1]].reshape(h, w, 4)
                ref = orc.slice_segment(rgba, comp, int(k), x0, y0, w, h)
                if swap:
                    ref = ref[..., [2, 1, 0, 3]]
                if not np.array_equal(sl, ref) or keys[v, k] != uniq[k]:
                    raise Mismatch(f"SLICE MISMATCH {recipe} map {v} segment {int(k)}")
            f0 += det["K"][v]
        if offh[-1] != bufh.size or len(offh) != f0 + 1:
            raise Mismatch(f"SLICE PLAN MISMATCH {recipe}")
        canvas = ws.reassemble_map(0, buf, off).cpu().numpy()
        want_canvas = orc.chw3_f32_to_rgba(images[0])
        if swap:
            want_canvas = want_canvas[..., [2, 1, 0, 3]]
        if not np.array_equal(canvas, want_canvas):
            raise Mismatch(f"DECODER MISMATCH {recipe}")
        ws.close()
    log(f"fuzz ok: {done} cases, seed {seed}, worst relative reward error {worst:.2e}, {time.time() - t0:.0f} s")
    return done, worst


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    try:
        run(iters, seed)
    except Mismatch as e:
        print(e)
        sys.exit(1)


if __name__ == "__main__":
    main()
