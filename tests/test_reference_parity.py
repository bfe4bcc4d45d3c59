"""GPU: the CUDA path against the reference's OWN CUDA sources run live on the same box
(oracle/_ref/libref_reward.so, built unchanged from /root/reference/src/*.cu by oracle/build_ref.sh),
and against the golden vectors that library produced (tests/golden/reference_cases.npz)."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "libref_reward.so")
GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_cases.npz")
RTOL = 1e-5


@pytest.fixture(scope="module")
def ref(dev):
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/libref_reward.so not built (needs /root/reference at build time)")
    L = C.CDLL(REF)
    L.ref_last_error.restype = C.c_char_p
    return L


def _golden():
    z = np.load(GOLDEN)
    names = sorted({k.split("/")[0] for k in z.files})
    return {n: {k.split("/")[1]: z[k] for k in z.files if k.startswith(n + "/")} for n in names}


@pytest.mark.parametrize("name", sorted(_golden()) if os.path.exists(GOLDEN) else [])
def test_cuda_path_matches_golden_reference_outputs(icr, dev, name):
    c = _golden()[name]
    p = c["params"]
    ti, ts = torch.from_numpy(c["images"]).to(dev), torch.from_numpy(c["sizes"]).to(dev)
    tl = torch.from_numpy(c["labels"]).to(dev)
    B, _, H, W = c["images"].shape
    ws = icr.Workspace(dev, B, H, W)
    r = icr.compute_rewards_batched(ti, tl, ts, int(p[0]), int(p[1]), float(p[2]), float(p[3]), float(p[4]), float(p[5]), bool(p[6]),
                                    float(p[7]), workspace=ws).cpu().numpy()
    assert np.allclose(r, c["rewards"], rtol=RTOL, atol=0)
    K, counts, bboxes, seg = [t.cpu().numpy() for t in ws.tables()]
    k = len(c["counts0"])
    assert K[0] == k
    assert np.array_equal(counts[0, :k], c["counts0"]) and np.array_equal(bboxes[0, :k], c["bboxes0"])
    assert np.allclose(seg[0, :k], c["seg_sizes0"], rtol=RTOL, atol=0)


def test_live_reference_rewards(icr, ref, dev):
    images, labels, sizes = synth.make_batch(2, B=5, seed0=777)
    ti, tl, ts = torch.from_numpy(images).to(dev), torch.from_numpy(labels).to(dev), torch.from_numpy(sizes).to(dev)
    out = torch.empty(5, dtype=torch.float64, device=dev)
    rc = ref.ref_compute_rewards(C.c_void_p(ti.data_ptr()), 1, C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), 5, 224, 224, 0,
                                 1, 4, C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1), C.c_double(9.308622), 1, C.c_double(0.5),
                                 C.c_void_p(out.data_ptr()))
    assert rc == 0, ref.ref_last_error()
    got = icr.compute_rewards_batched(ti, tl, ts)
    assert torch.allclose(got, out, rtol=RTOL, atol=0)


def test_compact_label_stats_dropin(icr, ref, orc, dev):
    lab = synth.bfs_partition(90, 130, 37, seed=9)
    comp, uniq = orc.compact_labels(lab)
    K = len(uniq)
    tl = torch.from_numpy(comp).to(dev)
    c0, b0 = torch.empty(K, dtype=torch.int32, device=dev), torch.empty((K, 4), dtype=torch.int32, device=dev)
    c1, b1 = torch.empty_like(c0), torch.empty_like(b0)
    assert ref.ref_segment_stats(C.c_void_p(tl.data_ptr()), 90, 130, K, 0, C.c_void_p(c0.data_ptr()), C.c_void_p(b0.data_ptr())) == 0
    icr.compute_counts_bboxes_from_compact_labels_cuda(tl, c1, b1)
    assert torch.equal(c0, c1) and torch.equal(b0, b1)


@pytest.mark.parametrize("channels", [1, 2, 3, 4])
@pytest.mark.parametrize("adaptive", [True, False])
def test_single_segment_estimator_dropin(icr, ref, orc, dev, channels, adaptive):
    H, W = 70, 90
    rng = np.random.default_rng(channels)
    img = np.ascontiguousarray(np.concatenate([synth.image_u8(3, H, W, 4), rng.integers(0, 256, (H, W, 1), dtype=np.uint8)], 2)[:, :, :channels])
    comp, uniq = orc.compact_labels(synth.bfs_partition(H, W, 6, seed=11))
    counts, bboxes = orc.segment_stats(comp, len(uniq))
    timg, tl, tc = torch.from_numpy(img).to(dev), torch.from_numpy(comp).to(dev), torch.from_numpy(counts).to(dev)
    for k in range(len(uniq)):
        x0, y0, x1, y1 = [int(v) for v in bboxes[k]]
        w, h = x1 - x0 + 1, y1 - y0 + 1
        want = torch.zeros(1, dtype=torch.float64, device=dev)
        rc = ref.ref_estimate_segment(C.c_void_p(timg.data_ptr()), W, H, channels, C.c_void_p(tl.data_ptr()), C.c_void_p(tc.data_ptr()),
                                      C.c_int64(k), x0, y0, w, h, 0, 1, 2, C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1),
                                      C.c_double(9.308622), int(adaptive), C.c_void_p(want.data_ptr()))
        assert rc == 0, ref.ref_last_error()
        got = icr.estimate_png_size_masked_segment_to_output(timg, tl, tc, k, x0, y0, w, h, L_min=2, adaptive_filter=adaptive)
        assert torch.allclose(got, want, rtol=RTOL, atol=0), (channels, k, got, want)
    # w <= 0 writes 0.0 (src/png_size_estimator.cu:485-488); channels outside 1..4 is an error (:484)
    z = icr.estimate_png_size_masked_segment_to_output(timg, tl, tc, 0, 0, 0, 0, 5, out=torch.ones(1, dtype=torch.float64, device=dev))
    assert z.item() == 0.0
    with pytest.raises(RuntimeError, match="channels must be 1..4"):
        icr.estimate_png_size_masked_segment_to_output(torch.zeros((4, 4, 5), dtype=torch.uint8, device=dev), tl, tc, 0, 0, 0, 2, 2)


@pytest.mark.parametrize("channels,H,W", [(1, 400, 420), (2, 300, 330), (3, 250, 300), (4, 300, 290)])
def test_single_segment_estimator_dropin_with_match_chunks(icr, ref, dev, channels, H, W):
    """Segments past 65536 stream bytes: chunk = 3, 4, 4, 6 bytes here, not a whole number of pixels, so Rule M's run
    statistics (src/png_size_estimator.cu:416-451) cross pixel and tile boundaries of the big-segment kernel."""
    rng = np.random.default_rng(10 + channels)
    img = np.ascontiguousarray(np.concatenate([synth.image_u8(5, H, W, 4), rng.integers(0, 256, (H, W, 1), dtype=np.uint8)], 2)[:, :, :channels])
    img[60:200, 40:240] = 93                                  # flat area: runs well past L_min
    lab = np.zeros((H, W), np.int64)
    lab[10:40, 10:50] = 1
    counts = np.bincount(lab.reshape(-1), minlength=2).astype(np.int32)
    timg, tl, tc = torch.from_numpy(img).to(dev), torch.from_numpy(lab).to(dev), torch.from_numpy(counts).to(dev)
    for L_min in (4, 2):
        want = torch.zeros(1, dtype=torch.float64, device=dev)
        rc = ref.ref_estimate_segment(C.c_void_p(timg.data_ptr()), W, H, channels, C.c_void_p(tl.data_ptr()), C.c_void_p(tc.data_ptr()),
                                      C.c_int64(0), 0, 0, W, H, 0, 1, L_min, C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1),
                                      C.c_double(9.308622), 1, C.c_void_p(want.data_ptr()))
        assert rc == 0, ref.ref_last_error()
        got = icr.estimate_png_size_masked_segment_to_output(timg, tl, tc, 0, 0, 0, W, H, L_min=L_min, adaptive_filter=True)
        assert torch.allclose(got, want, rtol=RTOL, atol=0), (channels, L_min, got, want)


def test_ema_baseline_and_advantage(icr, orc, dev):
    rng = np.random.default_rng(5)
    ema = icr.EMABaseline(0.99)
    val, init = 0.0, False
    for step in range(4):
        r = rng.normal(0.2, 0.3, 37)
        tr = torch.from_numpy(r).to(dev)
        v, adv = ema.update(tr, advantages=True)
        val, init = orc.ema_update(val, init, 0.99, r)
        assert abs(v.item() - val) < 1e-14
        a = r - val                                        # src/fcn/training.cpp:179-181
        a = (a - a.mean()) / max(a.std(), 1e-6)
        assert np.allclose(adv.cpu().numpy(), a, rtol=1e-10, atol=1e-12)
    sc = icr.reward_sum_count(tr).cpu().numpy()
    assert abs(sc[0] - r.sum()) < 1e-12 and sc[1] == 37
