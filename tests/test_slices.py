"""The rows after the reward path: all slices of a batch in one packed buffer,
the metadata.bin sidecar, and the decoder's paste.  CPU tests pin the restatement (oracle/slices_oracle.py) against the
bytes the reference's own writer produced; GPU tests compare the CUDA path, through the C ABI, with the restatement."""
import os
import sys

import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth
from oracle import slices_oracle as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, GOLDEN)


# ------------------------------------------------------------------------------------------------------------- CPU
def test_metadata_restatement_matches_the_reference_writer_bytes():
    import make_metadata_golden as g
    ref = open(os.path.join(GOLDEN, "metadata_ref.bin"), "rb").read()
    assert so.metadata_bytes(g.RECORDS, g.IMAGE_W, g.IMAGE_H) == ref
    recs, iw, ih = so.read_metadata(ref)
    assert (iw, ih) == (g.IMAGE_W, g.IMAGE_H) and recs == list(g.RECORDS)
    with pytest.raises(RuntimeError, match="magic"):
        so.read_metadata(b"\0" * 16)


def test_oracle_slices_round_trip_through_the_decoder(orc):
    """Size-independent property: the slices of a partition paste back to the source image."""
    H, W = 40, 56
    rgba = orc.chw3_f32_to_rgba(synth.image_f32_chw(2, H, W, 3))
    lab = synth.bfs_partition(H, W, 9, seed=2)
    recs = so.write_slices(rgba, lab)
    assert [r[0] for r in recs] == sorted(set(lab.reshape(-1).tolist()))
    comp, uniq = orc.compact_labels(lab)
    for r in recs:                                   # the batch oracle's Rule P and the slicer's restatement agree
        k = int(np.searchsorted(uniq, r[0]))
        assert np.array_equal(r[6], orc.slice_segment(rgba, comp, k, r[2], r[3], r[4], r[5]))
    meta, iw, ih = so.read_metadata(so.metadata_bytes(recs, W, H))
    assert np.array_equal(so.reassemble(meta, [r[6] for r in recs], iw, ih), rgba)
    # a slice hanging over the canvas edge is clipped, one outside is skipped (src/reassemble.cpp:58-64)
    sl = np.full((4, 4, 4), 255, np.uint8)
    c = so.reassemble([(0, "a", W - 2, H - 1, 4, 4), (1, "b", W, 0, 4, 4)], [sl, sl], W, H)
    assert c[..., 3].sum() == 255 * 2 and c[H - 1, W - 2:, 3].tolist() == [255, 255]


# ------------------------------------------------------------------------------------------------------------- GPU
def _batch(icr, dev, images, labels, sizes, prepare_only=False):
    import ctypes as C
    V, H, W = int(np.prod(labels.shape[:-2])), labels.shape[-2], labels.shape[-1]
    ws = icr.Workspace(dev, V, H, W, None)
    ti, tl, ts = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (images, labels, sizes))
    if prepare_only:
        S = labels.shape[1] if labels.ndim == 4 else 1
        code = 1 if labels.dtype == np.int64 else 0
        rc = icr.lib().icr_slices_prepare(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), code, images.shape[0], S, H, W,
                                          ws.handle, C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, icr.lib().icr_last_error()
        ws.check()
    else:
        icr.compute_rewards_batched(ti, tl, ts, workspace=ws)
    ws.inputs = (ti, tl, ts)
    return ws


def _compare_with_slicer(icr, orc, ws, images, labels, S=1):
    buf, offsets = ws.slices()
    swapped, offsets2 = ws.slices(order=1)
    assert torch.equal(offsets, offsets2)
    buf, swapped, off = buf.cpu().numpy(), swapped.cpu().numpy(), offsets.cpu().numpy()
    K, counts, bboxes, _ = [t.cpu().numpy() for t in ws.tables()]
    keys = ws.keys().cpu().numpy()
    maps = labels.reshape(-1, labels.shape[-2], labels.shape[-1])
    f = 0
    for v in range(maps.shape[0]):
        rgba = orc.chw3_f32_to_rgba(images[v // S])
        base = int(maps[v].min())
        recs = so.write_slices(rgba, maps[v] - base)            # the reference enumerates labels from 0: shift, then shift back
        assert K[v] == len(recs)
        for k, (label, _, x, y, w, h, sl) in enumerate(recs):
            assert keys[v, k] == label + base and bboxes[v, k].tolist() == [x, y, x + w - 1, y + h - 1]
            got = buf[off[f]:off[f + 1]].reshape(h, w, 4)
            assert np.array_equal(got, sl), f"slice bytes map {v} segment {k}"
            assert np.array_equal(swapped[off[f]:off[f + 1]].reshape(h, w, 4), sl[..., [2, 1, 0, 3]])
            f += 1
    assert f == len(off) - 1 and off[-1] == buf.size
    return buf, off


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_batch_slices_match_the_slicer_restatement(icr, orc, dev, dtype):
    images, labels, sizes = synth.make_batch(2, B=3)
    ws = _batch(icr, dev, images, labels.astype(dtype), sizes)
    _compare_with_slicer(icr, orc, ws, images, labels)


@pytest.mark.gpu
def test_slices_of_irregular_partitions_and_decoder_round_trip(icr, orc, dev):
    H, W = 72, 100                                              # bboxes overlap: total slice bytes exceed the image
    images = np.stack([synth.image_f32_chw(f, H, W, 10 + f) for f in range(3)])
    labels = np.stack([synth.bfs_partition(H, W, 11, seed=1), synth.voronoi_partition(H, W, cell=12, jitter=5, seed=2),
                       np.zeros((H, W), np.int64)]).astype(np.int32)          # the last map is one segment: the whole image
    ws = _batch(icr, dev, images, labels, np.full(3, 9000.0), prepare_only=True)
    buf, off = _compare_with_slicer(icr, orc, ws, images, labels)
    assert off[-1] > 3 * H * W * 4
    tb, to = torch.from_numpy(buf).to(dev), torch.from_numpy(off).to(dev)
    for v in range(3):
        canvas = ws.reassemble_map(v, tb, to).cpu().numpy()
        assert np.array_equal(canvas, orc.chw3_f32_to_rgba(images[v])), f"decoder round trip of map {v}"


@pytest.mark.gpu
def test_multi_map_slices_share_the_image(icr, orc, dev):
    images, labels, sizes = synth.make_multi_batch(B=2, S=3, H=48, W=64)
    ws = _batch(icr, dev, images, labels, sizes)
    _compare_with_slicer(icr, orc, ws, images, labels, S=3)


@pytest.mark.gpu
def test_paste_slice_is_sequential_and_clipped(icr, dev):
    """src/reassemble.cpp:58-97: later slices win where two overlap, transparent pixels leave the canvas alone."""
    H, W = 20, 30
    rng = np.random.default_rng(3)
    recs, slices = [], []
    for i, (x, y, w, h) in enumerate([(0, 0, 12, 9), (6, 4, 20, 10), (25, 15, 9, 9), (30, 0, 4, 4), (3, 19, 5, 1)]):
        sl = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        sl[..., 3] = np.where(rng.random((h, w)) < 0.4, 0, sl[..., 3] | 1)
        recs.append((i, "s%d" % i, x, y, w, h))
        slices.append(sl)
    canvas = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
    for r, sl in zip(recs, slices):
        icr.paste_slice(canvas, torch.from_numpy(sl).to(dev), r[2], r[3])
    assert np.array_equal(canvas.cpu().numpy(), so.reassemble(recs, slices, W, H))
    with pytest.raises(RuntimeError, match="geometry"):
        icr.paste_slice(canvas, torch.from_numpy(slices[0]).to(dev), -1, 0)


@pytest.mark.gpu
def test_slice_entries_report_misuse(icr, dev):
    import ctypes as C
    ws = icr.Workspace(dev, 2, 16, 16, None)
    off = torch.zeros(8, dtype=torch.int64, device=dev)
    L = icr.lib()
    assert L.icr_slices_plan(ws.handle, C.c_void_p(off.data_ptr()), C.c_longlong(8), C.c_void_p(off.data_ptr()), None) == 1
    assert b"no batch call" in L.icr_last_error()
    assert L.icr_slices_emit(ws.handle, C.c_void_p(off.data_ptr()), C.c_void_p(off.data_ptr()), C.c_size_t(64), 7, None) == 1
    assert b"order" in L.icr_last_error()


@pytest.mark.gpu
@pytest.mark.parametrize("H,W", [(1, 1), (7, 30), (19, 1), (40, 52)])
def test_slices_of_odd_shapes(icr, orc, dev, H, W):
    rng = np.random.default_rng(H * 100 + W)
    images = (rng.random((3, 3, H, W), dtype=np.float32) * 255.0).astype(np.float32)
    yy, xx = np.mgrid[0:H, 0:W]
    labels = np.stack([np.zeros((H, W), np.int64), (yy * W + xx) % 7, (xx // 4) + 3 * (yy // 5)]).astype(np.int32)
    ws = _batch(icr, dev, images, labels, np.full(3, 500.0), prepare_only=True)
    buf, off = _compare_with_slicer(icr, orc, ws, images, labels)
    tb, to = torch.from_numpy(buf).to(dev), torch.from_numpy(off).to(dev)
    batch = ws.reassemble_batch(tb, to).cpu().numpy()              # all maps in one launch: the same canvases
    for v in range(3):
        assert np.array_equal(ws.reassemble_map(v, tb, to).cpu().numpy(), orc.chw3_f32_to_rgba(images[v]))
        assert np.array_equal(batch[v], orc.chw3_f32_to_rgba(images[v]))
