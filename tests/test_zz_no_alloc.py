"""GPU: icr_estimate_png_size_segment_noalloc, the single-segment estimate with its filter-id scratch in a static device
array (no cudaMalloc of any kind; include/icr_b200.h).  Called through the C ABI, twice in a row, and compared with the
existing entry point icr_estimate_png_size_segment on the same inputs: same arithmetic, so the size is required bit for
bit (the suite's tolerance for sizes is 1e-5 relative, BASELINE.json north_star) and the debug tables exactly.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-5
ICR_ECAPACITY = 3
ICR_NOALLOC_MAX_ROWS = 16384


def _call(L, fn, timg, tl, tc, k, box, L_min, adaptive, out, tabs):
    H, W, ch = timg.shape
    x0, y0, w, h = box
    ptrs = [C.c_void_p(t.data_ptr()) for t in tabs] if tabs else [None] * 4
    return fn(C.c_void_p(timg.data_ptr()), C.c_int(W), C.c_int(H), C.c_int(ch), C.c_void_p(tl.data_ptr()), C.c_void_p(tc.data_ptr()),
              C.c_int64(k), C.c_int(x0), C.c_int(y0), C.c_int(w), C.c_int(h), C.c_int32(1), C.c_int(L_min), C.c_float(0.012167),
              C.c_float(18.0), C.c_float(0.1), C.c_double(9.308622), C.c_int(int(adaptive)), C.c_void_p(out.data_ptr()), *ptrs,
              C.c_void_p(torch.cuda.current_stream().cuda_stream))


def _tables(h, ch, dev):
    return (torch.zeros((h, 5), dtype=torch.int32, device=dev), torch.zeros(h, dtype=torch.uint8, device=dev),
            torch.zeros((ch, 256), dtype=torch.int32, device=dev), torch.zeros(3, dtype=torch.int64, device=dev))


@pytest.mark.parametrize("channels,H,W", [(4, 96, 120), (3, 250, 300), (4, 300, 290)])
def test_noalloc_entry_matches_allocating_entry_twice_in_a_row(icr, orc, dev, channels, H, W):
    L = icr.lib()
    rng = np.random.default_rng(40 + channels + H)
    img = np.ascontiguousarray(np.concatenate([synth.image_u8(5, H, W, 4), rng.integers(0, 256, (H, W, 1), dtype=np.uint8)], 2)[:, :, :channels])
    img[H // 4:H // 2, W // 8:W // 2] = 93                                    # flat area: match runs past L_min (Rule M) when the stream is long
    comp, uniq = orc.compact_labels(synth.bfs_partition(H, W, 5, seed=H))
    counts, bboxes = orc.segment_stats(comp, len(uniq))
    timg, tl, tc = torch.from_numpy(img).to(dev), torch.from_numpy(comp).to(dev), torch.from_numpy(counts).to(dev)
    cases = [(k, (int(b[0]), int(b[1]), int(b[2] - b[0] + 1), int(b[3] - b[1] + 1))) for k, b in enumerate(bboxes)]
    cases.append((0, (0, 0, W, H)))                                             # the whole image as one bbox: the tallest scratch use
    with torch.cuda.device(dev):
        for rep in range(2):                                                     # two calls in a row per case: the array keeps the last call's bytes
            for adaptive in (True, False):
                for k, box in cases:
                    h = box[3]
                    want, got = (torch.full((1,), -1.0, dtype=torch.float64, device=dev) for _ in range(2))
                    wt, gt = _tables(h, channels, dev), _tables(h, channels, dev)
                    assert _call(L, L.icr_estimate_png_size_segment, timg, tl, tc, k, box, 4, adaptive, want, wt) == 0, L.icr_last_error()
                    assert _call(L, L.icr_estimate_png_size_segment_noalloc, timg, tl, tc, k, box, 4, adaptive, got, gt) == 0, L.icr_last_error()
                    torch.cuda.synchronize()
                    assert want.item() > 0
                    assert torch.allclose(got, want, rtol=RTOL, atol=0), (rep, adaptive, k, got, want)
                    assert got.item() == want.item(), "same kernel, same arithmetic: expected bit for bit"
                    for a, b, name in zip(gt, wt, ("row costs", "filter ids", "histogram", "match stats")):
                        assert torch.equal(a, b), (name, rep, adaptive, k)
    # against the CPU oracle as well, so that both entries being wrong together would show (4 channels: Rule P bytes)
    if channels == 4:
        k, (x0, y0, w, h) = cases[1]
        got = torch.zeros(1, dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            assert _call(L, L.icr_estimate_png_size_segment_noalloc, timg, tl, tc, k, (x0, y0, w, h), 4, True, got, None) == 0
        ref = orc.estimate_segment(img, comp, k, x0, y0, w, h)
        assert abs(got.item() - ref) <= RTOL * abs(ref)


def test_noalloc_entry_edge_cases_and_limit(icr, dev):
    L = icr.lib()
    H, W = 8, 8
    timg = torch.zeros((H, W, 4), dtype=torch.uint8, device=dev)
    tl = torch.zeros((H, W), dtype=torch.int64, device=dev)
    tc = torch.full((1,), H * W, dtype=torch.int32, device=dev)
    out = torch.ones(1, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        # w <= 0 or h <= 0 writes 0.0, as the allocating entry does (src/png_size_estimator.cu:485-488)
        assert _call(L, L.icr_estimate_png_size_segment_noalloc, timg, tl, tc, 0, (0, 0, 0, 5), 4, True, out, None) == 0
        assert out.item() == 0.0
        # a bbox taller than the static scratch is refused with a message that names the limit, and nothing is launched
        tall = ICR_NOALLOC_MAX_ROWS + 1
        timg2 = torch.zeros((tall, 2, 1), dtype=torch.uint8, device=dev)
        tl2 = torch.zeros((tall, 2), dtype=torch.int64, device=dev)
        out.fill_(7.0)
        rc = _call(L, L.icr_estimate_png_size_segment_noalloc, timg2, tl2, tc, 0, (0, 0, 2, tall), 4, True, out, None)
        assert rc == ICR_ECAPACITY
        msg = L.icr_last_error().decode()
        assert str(ICR_NOALLOC_MAX_ROWS) in msg and "ICR_NOALLOC_MAX_ROWS" in msg
        assert out.item() == 7.0
        # exactly at the limit it runs, and agrees with the allocating entry
        timg3, tl3 = timg2[:ICR_NOALLOC_MAX_ROWS].contiguous(), tl2[:ICR_NOALLOC_MAX_ROWS].contiguous()
        timg3[::3] = 200
        a, b = (torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(2))
        assert _call(L, L.icr_estimate_png_size_segment, timg3, tl3, tc, 0, (0, 0, 2, ICR_NOALLOC_MAX_ROWS), 4, True, a, None) == 0
        assert _call(L, L.icr_estimate_png_size_segment_noalloc, timg3, tl3, tc, 0, (0, 0, 2, ICR_NOALLOC_MAX_ROWS), 4, True, b, None) == 0
        assert a.item() == b.item() and a.item() > 0
        # bad arguments are refused like the allocating entry refuses them
        assert _call(L, L.icr_estimate_png_size_segment_noalloc, timg, tl, tc, 0, (4, 4, 8, 8), 4, True, out, None) != 0
        assert "bbox outside the image" in L.icr_last_error().decode()
