"""The C++ host layer (image-compression-cpp_b200/host: the reference's call surface over libtorch tensors, on top of
the C ABI) driven through its extern "C" test hooks, which take the same argument lists as oracle/ref_shim.cpp."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_LIB = os.path.join(ROOT, "image-compression-cpp_b200", "libicr_host.so")
RTOL = 1e-5


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(HOST_LIB):
        pytest.fail(f"{HOST_LIB} is missing: run __graft_entry__.build()")
    L = C.CDLL(HOST_LIB)
    L.icrh_last_error.restype = C.c_char_p
    return L


def test_host_library_exports_the_reference_call_surface():
    """CPU check: the drop-in C++ symbols exist with the reference's exact (mangled) signatures."""
    assert os.path.exists(HOST_LIB), "libicr_host.so has not been built"
    out = subprocess.check_output(["nm", "-DC", "--defined-only", HOST_LIB], text=True)
    for sig in ("compute_rewards_batched(at::Tensor const&, at::Tensor const&, at::Tensor const&, int, int, float, float, float, double, bool, double)",
                "compute_counts_bboxes_from_compact_labels_cuda(at::Tensor const&, at::Tensor&, at::Tensor&)",
                "estimate_png_size_masked_segment_to_output(unsigned char const*, int, int, int, long const*, int const*, long, int, int, int, int, int, int, float, float, float, double, bool, double*)",
                "EMABaseline::update(at::Tensor const&)", "labels_from_rama(at::Tensor const&, long, long)"):
        assert sig in out, sig
    for hook in ("icrh_compute_rewards", "icrh_segment_stats", "icrh_estimate_segment", "icrh_ema_step", "icrh_labels_from_rama"):
        assert hook in out


def _dev(*arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


def _p(t):
    return C.c_void_p(t.data_ptr())


def _rewards(host, images, labels, sizes, images_on_device=True, prm=(1, 4, 0.012167, 18.0, 0.1, 9.308622, 1, 0.5)):
    B, _, H, W = images.shape
    tl, ts = _dev(labels, sizes)
    ti = torch.from_numpy(images).cuda() if images_on_device else torch.from_numpy(images)
    out = torch.zeros(B, dtype=torch.float64, device="cuda")
    rc = host.icrh_compute_rewards(_p(ti), int(images_on_device), _p(tl), int(labels.dtype == np.int64), _p(ts), B, H, W, 0,
                                   prm[0], prm[1], C.c_float(prm[2]), C.c_float(prm[3]), C.c_float(prm[4]), C.c_double(prm[5]), prm[6],
                                   C.c_double(prm[7]), _p(out))
    return rc, out.cpu().numpy()


@pytest.mark.gpu
def test_cpp_compute_rewards_matches_oracle(host, orc, dev):
    images, labels, sizes = synth.make_batch(2, B=6)
    rc, got = _rewards(host, images, labels, sizes)
    assert rc == 0, host.icrh_last_error()
    want = orc.compute_rewards(images, labels, sizes)
    assert np.max(np.abs(got - want) / np.abs(want)) < RTOL
    # int64 labels and host-resident images go through the same entry (src/compute_rewards.cu:82-85,96-97)
    rc, got64 = _rewards(host, images, labels.astype(np.int64), sizes, images_on_device=False)
    assert rc == 0, host.icrh_last_error()
    assert np.array_equal(got, got64)


@pytest.mark.gpu
def test_cpp_compute_rewards_any_integer_labels_fall_back_to_ranking(host, orc, dev):
    """The reference sorts, so any integer values are legal; values spread beyond the bitmap are ranked first."""
    images, labels, sizes = synth.make_batch(2, B=2)
    wide = labels.astype(np.int64) * (1 << 33) - (1 << 50)
    rc, got = _rewards(host, images, wide, sizes)
    assert rc == 0, host.icrh_last_error()
    want = orc.compute_rewards(images, labels, sizes)          # ranking preserves order, so the rewards are those of the small labels
    assert np.max(np.abs(got - want) / np.abs(want)) < RTOL


@pytest.mark.gpu
def test_cpp_compute_rewards_grows_tables_past_first_capacity(host, orc, dev):
    H = W = 96                                                   # 4608 segments > the 2048 rows the cache starts with
    images = np.stack([synth.image_f32_chw(5, H, W, 3)])
    labels = (np.arange(H * W, dtype=np.int32).reshape(1, H, W) // 2)
    rc, got = _rewards(host, images, labels, np.array([20000.0]))
    assert rc == 0, host.icrh_last_error()
    want = orc.compute_rewards(images, labels, np.array([20000.0]))
    assert abs(got[0] - want[0]) <= RTOL * abs(want[0])


@pytest.mark.gpu
def test_cpp_argument_errors_are_the_references(host, dev):
    images, labels, sizes = synth.make_batch(2, B=2)
    B, _, H, W = images.shape
    ti, tl, ts = _dev(images, labels, sizes)
    out = torch.zeros(B, dtype=torch.float64, device="cuda")
    counts = torch.zeros(4, dtype=torch.int32, device="cuda")
    bb = torch.zeros(4, 4, dtype=torch.int32, device="cuda")
    lab = torch.zeros(8, 8, dtype=torch.int64, device="cuda")
    rc = host.icrh_estimate_segment(None, 8, 8, 4, _p(lab), _p(counts), 0, 0, 0, 8, 8, 0, 1, 4, C.c_float(0), C.c_float(0), C.c_float(0),
                                    C.c_double(0), 1, _p(out))
    assert rc == 2 and b"null ptr" in host.icrh_last_error()      # std::runtime_error, src/png_size_estimator.cu:483
    img = torch.zeros(8, 8, 4, dtype=torch.uint8, device="cuda")
    rc = host.icrh_estimate_segment(_p(img), 8, 8, 5, _p(lab), _p(counts), 0, 0, 0, 8, 8, 0, 1, 4, C.c_float(0), C.c_float(0), C.c_float(0),
                                    C.c_double(0), 1, _p(out))
    assert rc == 2 and b"channels" in host.icrh_last_error()      # :484
    node = torch.zeros(2, 64, dtype=torch.int32, device="cuda")
    assert host.icrh_labels_from_rama(_p(node), 2, 8, 8, 0, 0) == 0
    assert host.icrh_labels_from_rama(_p(node), 2, 8, 8, 0, 1) == 1 and b"contiguous" in host.icrh_last_error()


@pytest.mark.gpu
def test_cpp_segment_stats_and_estimator_entries(host, orc, dev):
    H, W = 72, 88
    lab = synth.bfs_partition(H, W, 7, seed=4).astype(np.int64)
    comp, uniq = orc.compact_labels(lab)
    K = len(uniq)
    tl = torch.from_numpy(comp).cuda()
    counts = torch.zeros(K, dtype=torch.int32, device="cuda")
    bboxes = torch.zeros(K, 4, dtype=torch.int32, device="cuda")
    assert host.icrh_segment_stats(_p(tl), H, W, K, 0, _p(counts), _p(bboxes)) == 0, host.icrh_last_error()
    wc, wb = orc.segment_stats(comp, K)
    assert np.array_equal(counts.cpu().numpy(), wc) and np.array_equal(bboxes.cpu().numpy(), wb)
    rgba = np.concatenate([synth.image_u8(3, H, W, 5), np.full((H, W, 1), 255, np.uint8)], 2)
    ti = torch.from_numpy(rgba).cuda()
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    for k in range(K):
        x0, y0, x1, y1 = wb[k]
        w, h = x1 - x0 + 1, y1 - y0 + 1
        rc = host.icrh_estimate_segment(_p(ti), W, H, 4, _p(tl), _p(counts), k, int(x0), int(y0), int(w), int(h), 0, 1, 4,
                                        C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1), C.c_double(9.308622), 1, _p(out))
        assert rc == 0, host.icrh_last_error()
        want = orc.estimate_segment(rgba, comp, k, x0, y0, w, h)
        assert abs(out.item() - want) <= RTOL * abs(want)
    rc = host.icrh_estimate_segment(_p(ti), W, H, 4, _p(tl), _p(counts), 0, 0, 0, 0, 5, 0, 1, 4, C.c_float(0.012167), C.c_float(18.0),
                                    C.c_float(0.1), C.c_double(9.308622), 1, _p(out))
    assert rc == 0 and out.item() == 0.0                          # w <= 0 writes 0.0, src/png_size_estimator.cu:485-488


@pytest.mark.gpu
def test_cpp_multi_map_entry_and_ema(host, orc, dev):
    images, labels, sizes = synth.make_multi_batch(B=2, S=3, H=48, W=64)
    ti, tl, ts = _dev(images, labels, sizes)
    out = torch.zeros(2, 3, dtype=torch.float64, device="cuda")
    assert host.icrh_compute_rewards_multi(_p(ti), _p(tl), int(labels.dtype == np.int64), _p(ts), 2, 3, 48, 64, 0, _p(out)) == 0, \
        host.icrh_last_error()
    want = orc.compute_rewards(np.repeat(images, 3, axis=0), labels.reshape(6, 48, 64), np.repeat(sizes, 3))
    got = out.cpu().numpy().reshape(-1)
    assert np.max(np.abs(got - want) / np.abs(want)) < RTOL
    # Rule E through the C++ struct: first call copies the mean, the second blends
    state = (C.c_double * 2)(0.0, 0.0)
    adv = torch.zeros(6, dtype=torch.float64, device="cuda")
    assert host.icrh_ema_step(_p(out), 6, 0, C.c_double(0.99), state, _p(adv)) == 0, host.icrh_last_error()
    v1, init = orc.ema_update(0.0, False, 0.99, got)
    assert init and abs(state[0] - v1) < 1e-12 and state[1] == 1.0
    a = (got - v1)
    a = (a - a.mean()) / max(a.std(), 1e-6)                       # src/fcn/training.cpp:178-181
    assert np.allclose(adv.cpu().numpy(), a, rtol=1e-9, atol=1e-12)
    assert host.icrh_ema_step(_p(out), 6, 0, C.c_double(0.99), state, None) == 0
    v2, _ = orc.ema_update(v1, True, 0.99, got)
    assert abs(state[0] - v2) < 1e-12
