"""Seeded synthetic inputs for the reward path: images and label maps (numpy, CPU).

The reference has no seeded generators: its image families live in the dead, unseeded
src/pattern_generator.cpp:42-220 and its partitions in src/pattern_generator.cpp:222-288
(BFS region growing) and include/random_partition.hpp:21-144 (BSP rectangles).  These are
seeded restatements of the same families so that oracle, reference build and CUDA path all
read identical bytes.  Images are float32 planes holding the integer
byte values 0..255 (training-style [0,1] inputs collapse to bytes 0/1).
"""
import numpy as np

FAMILIES = ("tile_repeat", "monochrome", "low_variance", "low_frequency", "row_copies", "uniform_noise")


def _bicubic_up(seed_hwc, H, W):
    try:
        import cv2
        return cv2.resize(seed_hwc, (W, H), interpolation=cv2.INTER_CUBIC)
    except Exception:  # pragma: no cover - cv2 is in the image, scipy is the stand-in
        from scipy.ndimage import zoom
        z = zoom(seed_hwc.astype(np.float32), (H / seed_hwc.shape[0], W / seed_hwc.shape[1], 1), order=3)
        return np.clip(np.rint(z), 0, 255).astype(np.uint8)


def image_u8(family, H, W, seed):
    """One HxWx3 uint8 image of the given family (index or name)."""
    if isinstance(family, int):
        family = FAMILIES[family % len(FAMILIES)]
    rng = np.random.default_rng(seed)
    if family == "tile_repeat":            # pattern_generator.cpp:42-53, 8x8 tile
        t = rng.integers(0, 256, (8, 8, 3), dtype=np.uint8)
        return np.tile(t, ((H + 7) // 8, (W + 7) // 8, 1))[:H, :W].copy()
    if family == "monochrome":             # :55-70
        return np.broadcast_to(rng.integers(0, 256, 3, dtype=np.uint8), (H, W, 3)).copy()
    if family == "low_variance":           # :72-111, mean 50..200, sigma 2..8, truncation toward zero
        mean = rng.integers(50, 201, 3)
        sig = rng.integers(2, 9, 3)
        v = rng.normal(mean, sig, (H, W, 3)).astype(np.float32)
        return np.clip(v.astype(np.int32), 0, 255).astype(np.uint8)
    if family == "low_frequency":          # :113-165, 32x32 seed, sigma 2..20, bicubic upscale
        mean = rng.integers(50, 201, 3)
        sig = rng.integers(2, 21, 3)
        s = np.clip(rng.normal(mean, sig, (32, 32, 3)).astype(np.int32), 0, 255).astype(np.uint8)
        return _bicubic_up(s, H, W)
    if family == "row_copies":             # :167-176
        row = rng.integers(0, 256, (1, W, 3), dtype=np.uint8)
        return np.broadcast_to(row, (H, W, 3)).copy()
    if family == "uniform_noise":          # :178-214
        return rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    raise ValueError(family)


def image_f32_chw(family, H, W, seed):
    """[3,H,W] float32 holding 0..255, the layout compute_rewards_batched takes."""
    return np.ascontiguousarray(image_u8(family, H, W, seed).transpose(2, 0, 1)).astype(np.float32)


def grid_partition(H, W, cell_h, cell_w):
    """Regular grid of cells (BASELINE config 1)."""
    ys = np.arange(H)[:, None] // cell_h
    xs = np.arange(W)[None, :] // cell_w
    return (ys * ((W + cell_w - 1) // cell_w) + xs).astype(np.int32)


def rect_partition(H, W, min_h=8, min_w=8, split_prob=0.75, min_rect_count=64, seed=1):
    """BSP rectangle tiling id map, after include/random_partition.hpp:21-144 (id-map half only)."""
    rng = np.random.default_rng(seed)
    pending = [(0, 0, W, H)]
    rects = []
    while pending:
        need_more = len(pending) + len(rects) < min_rect_count
        if need_more:    # largest splittable pending rectangle (:67-87)
            best, best_area = -1, -1
            for i, (x0, y0, w, h) in enumerate(pending):
                if (w >= 2 * min_w or h >= 2 * min_h) and w * h > best_area:
                    best, best_area = i, w * h
            if best < 0:
                rects.extend(pending)
                break
            r = pending[best]
            pending[best] = pending[-1]
            pending.pop()
        else:
            r = pending.pop()
        x0, y0, w, h = r
        can_v, can_h = w >= 2 * min_w, h >= 2 * min_h
        split = (can_v or can_h) and (need_more or rng.random() < split_prob)
        if not split:
            rects.append(r)
            continue
        split_v = (rng.random() < w / (w + h)) if (can_v and can_h) else can_v   # :104-111
        if split_v:
            sx = int(rng.integers(x0 + min_w, x0 + w - min_w + 1))
            pending += [(x0, y0, sx - x0, h), (sx, y0, x0 + w - sx, h)]
        else:
            sy = int(rng.integers(y0 + min_h, y0 + h - min_h + 1))
            pending += [(x0, y0, w, sy - y0), (x0, sy, w, y0 + h - sy)]
    ids = np.empty((H, W), np.int32)
    for rid, (x0, y0, w, h) in enumerate(rects):
        ids[y0:y0 + h, x0:x0 + w] = rid
    return ids


def bfs_partition(H, W, num_segments, seed=1):
    """Multi-seed randomised region growing, after pattern_generator.cpp:222-288.

    Vectorised: each round every unlabelled pixel next to a labelled one adopts a random
    labelled 4-neighbour with probability 1/2, which gives the same kind of ragged, connected
    regions as the reference's shuffled queue expansion."""
    rng = np.random.default_rng(seed)
    lab = np.full((H, W), -1, np.int32)
    ys = rng.integers(0, H, num_segments)
    xs = rng.integers(0, W, num_segments)
    for i in range(num_segments):
        lab[ys[i], xs[i]] = i
    while (lab < 0).any():
        pad = np.pad(lab, 1, constant_values=-1)
        nb = np.stack([pad[:-2, 1:-1], pad[2:, 1:-1], pad[1:-1, :-2], pad[1:-1, 2:]])   # up, down, left, right
        pick = rng.integers(0, 4, (H, W))
        cand = np.take_along_axis(nb, pick[None], 0)[0]
        grow = (lab < 0) & (cand >= 0) & (rng.random((H, W)) < 0.5)
        lab[grow] = cand[grow]
    return lab


def voronoi_partition(H, W, cell=16, jitter=6, seed=1):
    """SLIC-style map: nearest jittered grid seed (stand-in for cv::ximgproc SLIC, src/slic_edge.cpp:51-56)."""
    rng = np.random.default_rng(seed)
    gy, gx = (H + cell - 1) // cell, (W + cell - 1) // cell
    cy = (np.arange(gy)[:, None] * cell + cell / 2 + rng.integers(-jitter, jitter + 1, (gy, gx))).astype(np.float32)
    cx = (np.arange(gx)[None, :] * cell + cell / 2 + rng.integers(-jitter, jitter + 1, (gy, gx))).astype(np.float32)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    by = np.clip(yy // cell, 0, gy - 1).astype(np.int32)
    bx = np.clip(xx // cell, 0, gx - 1).astype(np.int32)
    best = np.full((H, W), np.inf, np.float32)
    out = np.zeros((H, W), np.int32)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            ny, nx = np.clip(by + dy, 0, gy - 1), np.clip(bx + dx, 0, gx - 1)
            d = (yy - cy[ny, nx]) ** 2 + (xx - cx[ny, nx]) ** 2
            idx = ny * gx + nx
            better = (d < best) | ((d == best) & (idx < out))
            best = np.where(better, d, best)
            out = np.where(better, idx, out)
    return out.astype(np.int32)


def make_batch(config, B=None, seed0=None):
    """(images f32 [B,3,H,W], labels int32 [B,H,W], image_sizes f64 [B]) for a BASELINE config.

    config 1: one 256x256, 8x8 grid of 32x32 cells.   config 2: 64 x 224^2 BSP rectangles (7*id+3).
    config 3: 256 x 512^2 Voronoi ("SLIC-style").     config 5: 4096 x 224^2 as config 2.
    image_sizes stands in for the on-disk PNG size (include/fcn/EdgeDataset.h:116): H*W*3/2 bytes.
    """
    if config == 1:
        B, H, W, seed0 = B or 1, 256, 256, 100 if seed0 is None else seed0
        lab = lambda b: grid_partition(H, W, 32, 32)
    elif config in (2, 5):
        B, H, W = B or (64 if config == 2 else 4096), 224, 224
        seed0 = (1000 if config == 2 else 5000) if seed0 is None else seed0
        # split_prob 0.1 keeps K near the ~64 segments BASELINE.json names (0.75 gives ~260)
        lab = lambda b: 7 * rect_partition(H, W, split_prob=0.1, seed=seed0 + b) + 3
    elif config == 3:
        B, H, W, seed0 = B or 256, 512, 512, 3000 if seed0 is None else seed0
        lab = lambda b: voronoi_partition(H, W, 16, 6, seed=seed0 + b)
    else:
        raise ValueError("config 4 has its own shape: use make_multi_batch")
    images = np.stack([image_f32_chw(b % 6, H, W, seed0 + 1000 + b) for b in range(B)])
    labels = np.stack([lab(b) for b in range(B)]).astype(np.int32)
    sizes = np.full(B, H * W * 1.5, np.float64)
    return images, labels, sizes


def make_multi_batch(B=256, S=8, H=224, W=224, seed0=4000):
    """Config 4: S label maps per image: BSP rectangles for even s, region growing for odd s."""
    images = np.stack([image_f32_chw(b % 6, H, W, seed0 + 100000 + b) for b in range(B)])
    labels = np.empty((B, S, H, W), np.int32)
    for b in range(B):
        for s in range(S):
            sd = seed0 + S * b + s
            k = 16 + (sd * 37) % 113
            labels[b, s] = rect_partition(H, W, min_rect_count=k, seed=sd) if s % 2 == 0 else bfs_partition(H, W, k, seed=sd)
    return images, labels, np.full(B, H * W * 1.5, np.float64)
