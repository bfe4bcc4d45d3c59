"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same seeded
inputs.  Bit-exact for areas, bboxes, slice bytes, row costs, filter choices, histograms and match triples;
1e-5 relative (BASELINE.json north_star) for the double-precision size estimate and reward."""
import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5      # BASELINE.json: "within 1e-5 relative for the floating-point size estimate and reward"


def _to(dev, *arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in arrs]


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.0


def _run(icr, dev, images, labels, sizes, Kmax=None, **kw):
    B, H, W = labels.shape[0], labels.shape[-2], labels.shape[-1]
    V = int(np.prod(labels.shape[:-2]))
    ws = icr.Workspace(dev, V, H, W, Kmax)
    ti, tl, ts = _to(dev, images, labels, sizes)
    r = icr.compute_rewards_batched(ti, tl, ts, workspace=ws, **kw)
    ws.inputs = (ti, tl, ts)          # the debug/slice entries read the caller's buffers of the last call: keep them alive
    return r.cpu().numpy(), ws


def _check_batch(icr, orc, dev, images, labels, sizes, tables=True, n_debug=6, Kmax=None, **kw):
    prm = orc.params(kw.get("min_pixels_per_segment", 1), kw.get("L_min", 4), kw.get("beta", 0.012167),
                     kw.get("b_match_token", 18.0), kw.get("gamma", 0.1), kw.get("overhead_base", 9.308622),
                     kw.get("adaptive_filter", True), kw.get("lambda_", 0.5))
    got, ws = _run(icr, dev, images, labels, sizes, Kmax=Kmax, **kw)
    B, _, H, W = images.shape
    Ks = ws.Kmax
    want, det = orc.compute_rewards(images, labels, sizes, prm, details=True, Kstride=Ks)
    K, counts, bboxes, seg = [t.cpu().numpy() for t in ws.tables()]
    assert np.array_equal(K[:B], det["K"])
    for b in range(B):
        k = det["K"][b]
        assert np.array_equal(counts[b, :k], det["counts"][b, :k]), f"counts of image {b}"
        assert np.array_equal(bboxes[b, :k], det["bboxes"][b, :k]), f"bboxes of image {b}"
        assert _rel(seg[b, :k], det["seg_sizes"][b, :k]) < RTOL, f"segment sizes of image {b}"
    assert _rel(got, want) < RTOL
    if not tables:
        return got, ws
    # Rule L: keys = sorted unique labels, compact ids = their ranks; Rule C/P bytes are checked through the slices below
    keys = ws.keys().cpu().numpy()
    rng = np.random.default_rng(7)
    for b in range(min(B, 3)):
        rgba = orc.chw3_f32_to_rgba(images[b])
        comp, uniq = orc.compact_labels(labels[b])
        assert np.array_equal(keys[b, :len(uniq)], uniq), "segment keys (sorted unique labels)"
        for k in rng.choice(det["K"][b], size=min(n_debug, det["K"][b]), replace=False):
            x0, y0, x1, y1 = det["bboxes"][b, k]
            w, h = x1 - x0 + 1, y1 - y0 + 1
            s, tab = orc.estimate_segment(rgba, comp, int(k), x0, y0, w, h, prm, debug=True)
            rc, fi, hist, match, size = [t.cpu().numpy() for t in ws.debug_segment(b, int(k), h, icr.make_params(
                prm.min_pixels, prm.L_min, prm.beta, prm.b_match_token, prm.gamma, prm.overhead_base, prm.adaptive_filter, prm.lambda_))]
            if prm.adaptive_filter:
                assert np.array_equal(rc.view(np.uint32), tab["row_costs"]), f"row costs b={b} k={k}"
            assert np.array_equal(fi, tab["filter_id"]), f"filter ids b={b} k={k}"
            assert np.array_equal(hist.view(np.uint32), tab["hist"]), f"histogram b={b} k={k}"
            assert np.array_equal(match.view(np.uint64), tab["match"]), f"match triple b={b} k={k}"
            assert abs(size[0] - s) <= RTOL * abs(s)
            sl = ws.slice_segment(b, int(k), w, h).cpu().numpy()
            assert np.array_equal(sl, orc.slice_segment(rgba, comp, int(k), x0, y0, w, h)), f"slice bytes b={b} k={k}"
    return got, ws


def test_selftest_packed_arithmetic(icr, dev):
    assert icr.selftest() == 0


def test_segment_stats_entry_with_compact_ids(icr, orc, dev):
    labels = np.stack([synth.bfs_partition(60, 84, 9, seed=b) * 11 - 40 for b in range(3)]).astype(np.int64)
    K, counts, bboxes, comp = icr.segment_stats(torch.from_numpy(labels).to(dev), compact=True)
    for b in range(3):
        want, uniq = orc.compact_labels(labels[b])
        wc, wb = orc.segment_stats(want, len(uniq))
        assert K[b].item() == len(uniq)
        assert np.array_equal(comp[b].cpu().numpy(), want.astype(np.int32))
        assert np.array_equal(counts[b, :len(uniq)].cpu().numpy(), wc) and np.array_equal(bboxes[b, :len(uniq)].cpu().numpy(), wb)


def test_config2_rect_partitions(icr, orc, dev):
    images, labels, sizes = synth.make_batch(2, B=12)
    _check_batch(icr, orc, dev, images, labels, sizes, Kmax=512)


def test_config1_grid_256(icr, orc, dev):
    images, labels, sizes = synth.make_batch(1)
    for fam in range(6):
        images[0] = synth.image_f32_chw(fam, 256, 256, 11 + fam)
        _check_batch(icr, orc, dev, images, labels, sizes, Kmax=64, n_debug=3)


def test_fine_rect_partitions_many_small_segments(icr, orc, dev):
    H = W = 96
    images = np.stack([synth.image_f32_chw(b, H, W, 50 + b) for b in range(6)])
    labels = np.stack([synth.rect_partition(H, W, split_prob=0.9, min_rect_count=20, seed=60 + b) * 5 - 17 for b in range(6)]).astype(np.int32)
    _check_batch(icr, orc, dev, images, labels, np.full(6, 9000.0))


def test_bfs_and_voronoi_partitions(icr, orc, dev):
    H, W = 120, 136
    images = np.stack([synth.image_f32_chw(b + 2, H, W, 70 + b) for b in range(4)])
    labels = np.stack([synth.bfs_partition(H, W, 23, seed=1), synth.voronoi_partition(H, W, 16, 6, seed=2),
                       synth.bfs_partition(H, W, 5, seed=3), synth.voronoi_partition(H, W, 9, 3, seed=4)]).astype(np.int32)
    _check_batch(icr, orc, dev, images, labels, np.full(4, 20000.0))


def test_int64_labels_and_negative_values(icr, orc, dev):
    images, labels, sizes = synth.make_batch(2, B=3)
    labels = labels.astype(np.int64) * 3 - 1000
    _check_batch(icr, orc, dev, images, labels, sizes, n_debug=2)


def test_odd_width_scalar_path(icr, orc, dev):
    H, W = 37, 53          # W % 4 != 0: the unvectorised prep path
    images = np.stack([synth.image_f32_chw(b, H, W, 90 + b) for b in range(3)])
    labels = np.stack([synth.bfs_partition(H, W, 7, seed=5 + b) for b in range(3)]).astype(np.int32)
    _check_batch(icr, orc, dev, images, labels, np.full(3, 3000.0))


def test_quirk_unit_range_images_collapse_to_bits(icr, orc, dev):
    rng = np.random.default_rng(3)
    images = rng.random((2, 3, 64, 64), dtype=np.float32)          # [0,1] training-style input
    labels = np.stack([synth.grid_partition(64, 64, 16, 16)] * 2).astype(np.int32)
    _check_batch(icr, orc, dev, images, labels, np.full(2, 5000.0))


def test_quirk_rounding_and_clamping(icr, orc, dev):
    rng = np.random.default_rng(4)
    base = rng.integers(-20, 280, (2, 3, 32, 32)).astype(np.float32)
    images = base + np.float32(0.5)                                   # x.5 everywhere: round half to even
    images[1] = base[1] + rng.choice(np.array([-0.5, 0.49, 0.51, 0.0], np.float32), base[1].shape)
    labels = np.stack([synth.grid_partition(32, 32, 8, 16)] * 2).astype(np.int32)
    _check_batch(icr, orc, dev, images, labels, np.full(2, 2000.0))


def test_quirk_single_segment_penalty(icr, orc, dev):
    images = np.stack([synth.image_f32_chw(2, 48, 48, 5)])
    labels = np.full((1, 48, 48), 9, np.int32)
    got, _ = _check_batch(icr, orc, dev, images, labels, np.array([4000.0]))
    # exactly one valid segment: the lambda penalty applies (src/compute_rewards.cu:183-192)
    got0, _ = _check_batch(icr, orc, dev, images, labels, np.array([4000.0]), lambda_=0.0)
    assert abs((got0[0] - got[0]) - 0.5) < 1e-12


def test_quirk_min_pixels_changes_only_the_penalty(icr, orc, dev):
    H = W = 40
    images = np.stack([synth.image_f32_chw(5, H, W, 6)])
    labels = np.zeros((1, H, W), np.int32)
    labels[0, 3, 4] = 1            # 1-pixel segments
    labels[0, 20, 21] = 2
    a, wa = _check_batch(icr, orc, dev, images, labels, np.array([3000.0]), min_pixels_per_segment=1)
    b, wb = _check_batch(icr, orc, dev, images, labels, np.array([3000.0]), min_pixels_per_segment=8)
    sa, sb = wa.tables()[3].cpu().numpy()[0, :3], wb.tables()[3].cpu().numpy()[0, :3]
    assert np.array_equal(sa, sb)                  # sizes untouched (skip_check_kernel is overwritten)
    assert abs((a[0] - b[0]) - 0.5) < 1e-12        # only the penalty term moved: 3 valid -> 1 valid


def test_quirk_non_adaptive_is_all_paeth(icr, orc, dev):
    images, labels, sizes = synth.make_batch(2, B=2)
    _check_batch(icr, orc, dev, images, labels, sizes, adaptive_filter=False, n_debug=3)


@pytest.mark.parametrize("L_min", [4, 2, 1, 7])
def test_rule_m_whole_image_bbox_224(icr, orc, dev, L_min):
    # one segment whose bbox is the whole 224x224 image: chunk = 4 bytes -> matches are possible at L_min<=4
    H = W = 224
    images = np.stack([synth.image_f32_chw(f, H, W, 8 + f) for f in (1, 2, 3)])
    labels = np.zeros((3, H, W), np.int32)
    labels[:, 60:90, 100:140] = 1
    labels[2] = synth.bfs_partition(H, W, 3, seed=2)
    _check_batch(icr, orc, dev, images, labels, np.full(3, 70000.0), L_min=L_min, n_debug=3)


def test_rule_m_whole_image_bbox_512(icr, orc, dev):
    H = W = 512               # chunk = 16 bytes = 4 pixels
    images = np.stack([synth.image_f32_chw(f, H, W, 8 + f) for f in (0, 3)])
    labels = np.zeros((2, H, W), np.int32)
    labels[:, 200:300, 100:400] = 5
    _check_batch(icr, orc, dev, images, labels, np.full(2, 300000.0), n_debug=2)


@pytest.mark.parametrize("H,W", [(256, 300), (300, 300), (320, 340)])
def test_rule_m_chunks_that_straddle_pixels(icr, orc, dev, H, W):
    # whole-image bboxes of 76800 / 90000 / 108800 pixels: chunk = 5, 6, 7 bytes, so chunks cut through pixels and the
    # stream tiles of the big-segment kernel end on a partial tile
    images = np.stack([synth.image_f32_chw(f, H, W, 20 + f) for f in (1, 4)])
    images[1, :, 40:200, 30:250] = 77.0                     # a flat area: long runs, matches in every chunk there
    labels = np.zeros((2, H, W), np.int32)
    labels[0, 100:140, 50:90] = 2
    _check_batch(icr, orc, dev, images, labels, np.full(2, 250000.0), n_debug=2)


def test_multi_label_maps_share_one_image(icr, orc, dev):
    images, labels, sizes = synth.make_multi_batch(B=3, S=4, H=64, W=80)
    ti, tl, ts = _to(dev, images, labels, sizes)
    got = icr.compute_rewards_batched(ti, tl, ts).cpu().numpy()
    assert got.shape == (3, 4)
    want = orc.compute_rewards(np.repeat(images, 4, axis=0), labels.reshape(12, 64, 80), np.repeat(sizes, 4))
    assert _rel(got.reshape(-1), want) < RTOL


def test_many_segments_global_stats_path(icr, orc, dev):
    # K > 2048 segments in one map: the shared-memory stats table is bypassed
    H = W = 128
    images = np.stack([synth.image_f32_chw(5, H, W, 3)])
    labels = (np.arange(H * W, dtype=np.int32).reshape(1, H, W) // 3) * 2
    _check_batch(icr, orc, dev, images, labels, np.array([30000.0]), n_debug=4)


def test_label_spread_beyond_capacity_is_reported(icr, dev):
    images, labels, sizes = synth.make_batch(2, B=2)
    labels = labels.astype(np.int64)
    labels[1, 0, 0] = 1 << 40
    ti, tl, ts = _to(dev, images, labels, sizes)
    with pytest.raises(RuntimeError, match="label"):
        icr.compute_rewards_batched(ti, tl, ts)
    r = icr.compute_rewards_batched(ti, tl, ts, check=False).cpu().numpy()
    assert np.isfinite(r[0]) and np.isnan(r[1])
    with pytest.raises(RuntimeError):          # the sticky flag is reported once, then cleared
        icr.get_workspace(dev, 2, 224, 224).check()
    icr.get_workspace(dev, 2, 224, 224).check()


def test_reference_argument_errors(icr, dev):
    images, labels, sizes = synth.make_batch(2, B=2)
    ti, tl, ts = _to(dev, images, labels, sizes)
    with pytest.raises(RuntimeError, match=r"images must be \[B,3,H,W\]"):
        icr.compute_rewards_batched(ti[:, :2], tl, ts)
    with pytest.raises(RuntimeError, match="float32"):
        icr.compute_rewards_batched(ti.double(), tl, ts)
    with pytest.raises(RuntimeError, match="labels must be CUDA"):
        icr.compute_rewards_batched(ti, tl.cpu(), ts)
    with pytest.raises(RuntimeError, match="image_sizes must be CUDA"):
        icr.compute_rewards_batched(ti, tl, ts.cpu())
    with pytest.raises(RuntimeError, match="B elements"):
        icr.compute_rewards_batched(ti, tl, ts[:1])
    with pytest.raises(RuntimeError, match="match images"):
        icr.compute_rewards_batched(ti, tl[:, :100], ts)
    # CPU images are moved to the labels' device, as the reference does (src/compute_rewards.cu:82-85)
    r = icr.compute_rewards_batched(ti.cpu(), tl, ts)
    assert r.is_cuda and r.dtype == torch.float64 and r.shape == (2,)


def test_full_size_config5_shard_properties(icr, orc, dev):
    # BASELINE config 5 shard (512 images of 224^2): size-independent properties + spot parity on a sample
    B = 512
    images, labels, sizes = synth.make_batch(5, B=64)
    reps = B // 64
    images, labels, sizes = np.tile(images, (reps, 1, 1, 1)), np.tile(labels, (reps, 1, 1)), np.tile(sizes, reps)
    got, ws = _run(icr, dev, images, labels, sizes, Kmax=256)
    K, counts, bboxes, seg = [t.cpu().numpy() for t in ws.tables()]
    assert np.all(counts[np.arange(256)[None, :] < K[:, None]] > 0)
    assert np.array_equal(np.where(np.arange(256)[None, :] < K[:, None], counts, 0).sum(1), np.full(B, 224 * 224))   # areas tile the image
    assert np.array_equal(got[:64], got[64:128]) and np.array_equal(got[:64], got[-64:])        # deterministic across the batch
    want = orc.compute_rewards(images[:8], labels[:8], sizes[:8])
    assert _rel(got[:8], want) < RTOL


@pytest.mark.parametrize("B,S", [(5, 1), (70, 1), (33, 2)])
def test_host_buffer_entry_matches_device_entry(icr, orc, dev, B, S):
    """icr_compute_rewards_host (what bench.py times as e2e): pageable host buffers in, rewards out; batches of 64 and more
    go over in four overlapped pieces with uneven sizes, and the result must not depend on that."""
    import ctypes as C
    H, W = 48, 64
    if S == 1:
        images, labels, sizes = synth.make_batch(2, B=min(B, 8))
        images, labels = images[:, :, :H, :W].copy(), labels[:, :H, :W].copy()
    else:
        images, labels, sizes = synth.make_multi_batch(B=8, S=S, H=H, W=W)
    reps = -(-B // images.shape[0])
    images = np.ascontiguousarray(np.tile(images, (reps, 1, 1, 1))[:B])
    labels = np.ascontiguousarray(np.tile(labels, (reps,) + (1,) * (labels.ndim - 1))[:B])
    sizes = np.ascontiguousarray(np.tile(sizes, reps)[:B])
    ws = icr.Workspace(dev, B * S, H, W, None)
    out = np.zeros(B * S, np.float64)
    prm = icr.make_params()
    L = icr.lib()
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    with torch.cuda.device(dev):
        rc = L.icr_compute_rewards_host(p(images), p(labels), int(labels.dtype == np.int64), p(sizes), B, S, H, W, C.byref(prm), p(out),
                                        ws.handle, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, L.icr_last_error()
    ws.check()
    ti, tl, ts = _to(dev, images, labels, sizes)
    want = icr.compute_rewards_batched(ti, tl, ts).cpu().numpy().reshape(-1)
    assert np.array_equal(out, want)
    n = min(B, 8)
    ref = orc.compute_rewards(np.repeat(images[:n], S, axis=0), labels[:n].reshape(n * S, H, W), np.repeat(sizes[:n], S))
    assert _rel(out[:n * S], ref) < RTOL
    if B >= 64:            # the tables hold the last piece only: the slice entries say so instead of returning part of the batch
        with pytest.raises(RuntimeError, match="no batch call"):
            ws.slices()


@pytest.mark.parametrize("B,S,layout,lab", [(5, 1, "planar", "u16"), (70, 1, "hwc", "u16"), (66, 2, "planar", "i32"), (9, 1, "hwc", "i64")])
def test_host_u8_entry_matches_f32_entry(icr, orc, dev, B, S, layout, lab):
    """icr_compute_rewards_host_u8: 8-bit pixels (planar or interleaved, what include/fcn/EdgeDataset.h:73 holds before its
    float conversion) and 16-bit label ids over the link, widened on the device.  (float)u8 is exact and Rule C maps it back
    to the byte, so the rewards must equal the f32 entry's on the converted tensors bit for bit; odd H*W exercises the
    scalar widening path, the larger batches the four overlapped pieces."""
    import ctypes as C
    H, W = (45, 63) if B < 10 else (48, 64)
    if S == 1:
        images, labels, sizes = synth.make_batch(2, B=min(B, 8))
        images, labels = images[:, :, :H, :W].copy(), labels[:, :H, :W].copy()
    else:
        images, labels, sizes = synth.make_multi_batch(B=8, S=S, H=H, W=W)
    reps = -(-B // images.shape[0])
    images = np.ascontiguousarray(np.tile(images, (reps, 1, 1, 1))[:B])
    labels = np.ascontiguousarray(np.tile(labels, (reps,) + (1,) * (labels.ndim - 1))[:B])
    sizes = np.ascontiguousarray(np.tile(sizes, reps)[:B])
    u8 = np.clip(np.rint(images), 0, 255).astype(np.uint8)                     # the bytes Rule C makes of the synthetic planes
    f32 = u8.astype(np.float32)
    host_img = np.ascontiguousarray(u8.transpose(0, 2, 3, 1)) if layout == "hwc" else u8
    if lab == "u16":
        assert labels.min() >= 0 and labels.max() < 65536
        host_lab, code = labels.astype(np.uint16), 2                              # ICR_LABELS_U16
    elif lab == "i64":
        host_lab, code = labels.astype(np.int64) - 7, 1
    else:
        host_lab, code = labels, 0
    ws = icr.Workspace(dev, B * S, H, W, None)
    out, want = np.zeros(B * S, np.float64), np.zeros(B * S, np.float64)
    prm = icr.make_params()
    L = icr.lib()
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    with torch.cuda.device(dev):
        for _ in range(2):                                                        # the second call reuses the staging buffers
            out[:] = 0
            rc = L.icr_compute_rewards_host_u8(p(host_img), int(layout == "hwc"), p(host_lab), code, p(sizes), B, S, H, W, C.byref(prm),
                                               p(out), ws.handle, st)
            assert rc == 0, L.icr_last_error()
            ws.check()
        rc = L.icr_compute_rewards_host(p(f32), p(labels), 0, p(sizes), B, S, H, W, C.byref(prm), p(want), ws.handle, st)
        assert rc == 0, L.icr_last_error()
        assert np.array_equal(out, want)
        # u16 labels only through this entry, and only the two layouts
        assert L.icr_compute_rewards_host(p(f32), p(labels), 2, p(sizes), B, S, H, W, C.byref(prm), p(want), ws.handle, st) != 0
        assert L.icr_compute_rewards_host_u8(p(host_img), 5, p(host_lab), code, p(sizes), B, S, H, W, C.byref(prm), p(out), ws.handle, st) != 0
    n = min(B, 8)
    ref = orc.compute_rewards(np.repeat(f32[:n], S, axis=0), labels[:n].reshape(n * S, H, W), np.repeat(sizes[:n], S))
    assert _rel(out[:n * S], ref) < RTOL


@pytest.mark.parametrize("H,W", [(1, 1), (1, 17), (17, 1), (3, 4), (4, 4), (2, 8), (8, 16), (5, 1000), (300, 4), (33, 36)])
def test_degenerate_shapes(icr, orc, dev, H, W):
    """One-pixel maps, single rows and columns, maps narrower than a quad and wider than a strip walk: every shape the
    reference accepts.  Label layouts: one segment, one segment per pixel, vertical and horizontal stripes."""
    rng = np.random.default_rng(H * 1000 + W)
    images = (rng.random((4, 3, H, W), dtype=np.float32) * 255.0).astype(np.float32)
    yy, xx = np.mgrid[0:H, 0:W]
    labels = np.stack([np.zeros((H, W), np.int64), (yy * W + xx).astype(np.int64), (xx // 3).astype(np.int64) * 5 - 7,
                       (yy // 2).astype(np.int64) + 100]).astype(np.int32)
    sizes = np.full(4, 64.0 + H * W)
    _check_batch(icr, orc, dev, images, labels, sizes, n_debug=4)
    _check_batch(icr, orc, dev, images, labels.astype(np.int64), sizes, tables=False)


def test_map_larger_than_the_label_bitmap(icr, orc, dev):
    """A 1024x768 map (786 432 pixels) is larger than the 2^19-bit presence bitmap shared memory holds: it works as long as
    the labels spread over fewer values than that, which any partition with fewer than 2^19 segments can be brought to."""
    H, W = 768, 1024
    images = np.stack([synth.image_f32_chw(1, H, W, 77)])
    labels = synth.voronoi_partition(H, W, cell=56, jitter=20, seed=9).astype(np.int32)[None] * 3 + 11
    got, ws = _check_batch(icr, orc, dev, images, labels, np.array([400000.0]), n_debug=3)
    assert icr.lib().icr_label_capacity(H, W) == 1 << 19
    # labels spread beyond the bitmap are reported, not mis-ranked
    wide = labels.astype(np.int64)
    wide[0, -1, -1] += 1 << 20
    ti, tl, ts = _to(dev, images, wide, np.array([400000.0]))
    with pytest.raises(RuntimeError, match="label spread"):
        icr.compute_rewards_batched(ti, tl, ts, workspace=ws)
