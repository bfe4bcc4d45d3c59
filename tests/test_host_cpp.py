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


# ---- the rows after the reward path: write_slices + metadata.bin + reassemble
def test_host_library_exports_the_slice_and_decoder_surface():
    out = subprocess.check_output(["nm", "-DC", "--defined-only", HOST_LIB], text=True)
    for sig in ("write_metadata_binary(std::vector<SliceMetadata, std::allocator<SliceMetadata> > const&, std::__cxx11::basic_string<char, std::char_traits<char>, std::allocator<char> > const&, unsigned int, unsigned int)",
                "read_metadata_binary(std::__cxx11::basic_string<char, std::char_traits<char>, std::allocator<char> > const&, unsigned int&, unsigned int&)",
                "write_slices(at::Tensor const&, at::Tensor const&, std::filesystem::__cxx11::path const&, std::filesystem::__cxx11::path const&)",
                "reassemble(std::filesystem::__cxx11::path, std::filesystem::__cxx11::path)",
                "slice_images_batched(at::Tensor const&, at::Tensor const&, bool)"):
        assert sig in out, sig


def test_cpp_metadata_writer_produces_the_reference_bytes(host, tmp_path):
    """No GPU involved: the C++ writer against the bytes src/metadata.cpp itself wrote (tests/golden/metadata_ref.bin)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_metadata_golden as g
    cols = [np.array([r[i] for r in g.RECORDS], np.int32) for i in (0, 2, 3, 4, 5)]
    names = b"".join(r[1].encode() + b"\0" for r in g.RECORDS)
    path = str(tmp_path / "metadata.bin").encode()
    assert host.icrh_write_metadata(path, len(g.RECORDS), *[c.ctypes.data_as(C.c_void_p) for c in cols], names,
                                    C.c_uint32(g.IMAGE_W), C.c_uint32(g.IMAGE_H)) == 0, host.icrh_last_error()
    assert open(path, "rb").read() == open(os.path.join(ROOT, "tests", "golden", "metadata_ref.bin"), "rb").read()
    iw, ih = C.c_uint32(), C.c_uint32()
    assert host.icrh_read_metadata_count(path, C.byref(iw), C.byref(ih)) == len(g.RECORDS) and (iw.value, ih.value) == (g.IMAGE_W, g.IMAGE_H)
    bad = tmp_path / "bad.bin"
    bad.write_bytes(b"\0" * 32)
    assert host.icrh_read_metadata_count(str(bad).encode(), C.byref(iw), C.byref(ih)) == -2 and b"magic mismatch" in host.icrh_last_error()
    assert host.icrh_read_metadata_count(str(tmp_path / "none.bin").encode(), C.byref(iw), C.byref(ih)) == -2
    assert b"Cannot open" in host.icrh_last_error()


@pytest.mark.gpu
def test_cpp_write_slices_files_and_reassemble_round_trip(host, orc, dev, tmp_path):
    from PIL import Image
    from oracle import slices_oracle as so
    H, W = 64, 88
    image = synth.image_f32_chw(1, H, W, 21)
    mask = synth.bfs_partition(H, W, 10, seed=6).astype(np.int32)
    ti, tm = _dev(image, mask)
    assert host.icrh_write_slices(_p(ti), _p(tm), 0, H, W, 0, str(tmp_path).encode(), b"img0") == 0, host.icrh_last_error()
    d = tmp_path / "img0"
    rgba = orc.chw3_f32_to_rgba(image)
    recs = so.write_slices(rgba, mask)
    assert (d / "metadata.bin").read_bytes() == so.metadata_bytes(recs, W, H)             # ascending labels, slice_<label>.png
    for label, name, x, y, w, h, sl in recs:
        assert np.array_equal(np.asarray(Image.open(d / name).convert("RGBA")), sl), name
    canvas = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda")
    out_png = tmp_path / "reconstructed.png"
    assert host.icrh_reassemble(str(d).encode(), str(out_png).encode(), 0, _p(canvas), H, W) == 0, host.icrh_last_error()
    assert np.array_equal(canvas.cpu().numpy(), rgba)
    assert np.array_equal(np.asarray(Image.open(out_png).convert("RGBA")), rgba)
    # the decoder reads files other encoders wrote too (RGB without alpha, every PNG filter type)
    x, y, w, h = recs[0][2:6]
    Image.fromarray(rgba[y:y + h, x:x + w, :3]).save(d / recs[0][1], optimize=True)
    assert host.icrh_reassemble(str(d).encode(), None, 0, _p(canvas), H, W) == 0, host.icrh_last_error()
    # that file is now opaque everywhere: its whole bbox is pasted first, the other slices land on top of it
    assert np.array_equal(canvas.cpu().numpy(), rgba)
    assert host.icrh_reassemble(str(tmp_path / "nothing").encode(), None, 0, _p(canvas), H, W) == 2
    assert b"Cannot open metadata file" in host.icrh_last_error()


@pytest.mark.gpu
def test_cpp_slice_image_by_label_value(host, orc, dev):
    """slice_image(input, mask, label, out_box) of include/image_slicer.h:6 / src/image_slicer.cpp:48-74 for one label VALUE:
    bbox + masked crop, bit for bit against the slicer restatement; a label that does not occur gives the empty result."""
    from oracle import slices_oracle as so
    H, W = 72, 100
    image = synth.image_f32_chw(3, H, W, 31)
    mask = (synth.bfs_partition(H, W, 9, seed=4) * 13 - 20).astype(np.int64)
    ti, tm = _dev(image, mask)
    rgba = orc.chw3_f32_to_rgba(image)
    out = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
    box = np.zeros(4, np.int32)
    for label in np.unique(mask)[[0, 3, -1]]:
        rc = host.icrh_slice_image(_p(ti), _p(tm), 1, H, W, 0, C.c_longlong(int(label)), box.ctypes.data_as(C.c_void_p), _p(out),
                                   C.c_longlong(out.numel()))
        assert rc == 0, host.icrh_last_error()
        sl, (x, y, w, h) = so.slice_image(rgba, mask, int(label))
        assert box.tolist() == [x, y, w, h]
        assert np.array_equal(out[:w * h * 4].cpu().numpy().reshape(h, w, 4), sl)
    rc = host.icrh_slice_image(_p(ti), _p(tm), 1, H, W, 0, C.c_longlong(12345), box.ctypes.data_as(C.c_void_p), _p(out),
                               C.c_longlong(out.numel()))
    assert rc == -1


@pytest.mark.gpu
def test_cpp_ema_value_is_a_fresh_tensor_per_update(host, dev):
    """include/ema_baseline.hpp:15-19 returns a new tensor each step: last step's baseline must not change under a caller
    that kept it."""
    r1, r2 = np.full(8, 2.0), np.full(8, 10.0)
    t1, t2 = _dev(r1, r2)
    vals = np.zeros(2)
    assert host.icrh_ema_two_steps(_p(t1), _p(t2), 8, 0, vals.ctypes.data_as(C.c_void_p)) == 0, host.icrh_last_error()
    assert vals[0] == 2.0 and abs(vals[1] - (0.9 * 2.0 + 0.1 * 10.0)) < 1e-12


@pytest.mark.gpu
def test_cpp_slices_with_labels_beyond_the_bitmap_are_ranked(host, orc, dev, tmp_path):
    """write_slices on a mask whose label spread defeats the presence bitmap: ranked on the fly, file names and sidecar
    still carry the caller's label values (int32 range, as the reference's SliceMetadata::label)."""
    from oracle import slices_oracle as so
    H, W = 40, 48
    image = synth.image_f32_chw(2, H, W, 5)
    mask = synth.grid_partition(H, W, 20, 16).astype(np.int64)
    mask[mask == mask.max()] = (1 << 30) + 7
    ti, tm = _dev(image, mask)
    assert host.icrh_write_slices(_p(ti), _p(tm), 1, H, W, 0, str(tmp_path).encode(), b"wide") == 0, host.icrh_last_error()
    rgba = orc.chw3_f32_to_rgba(image)
    recs = []                       # what src/image_slicer.cpp:96-117 records: the labels that occur, ascending
    for label in np.unique(mask):
        sl, (x, y, w, h) = so.slice_image(rgba, mask, int(label))
        recs.append((int(label), "slice_%d.png" % label, x, y, w, h, sl))
    assert (tmp_path / "wide" / "metadata.bin").read_bytes() == so.metadata_bytes(recs, W, H)


@pytest.mark.gpu
def test_node_mapping_packer_c_abi_and_cpp(host, icr, dev):
    """The producer side of the label handoff (src/rama_wrapper.cu:102-124): B per-sample node_mapping vectors -> one
    [B,H,W] int32 plane.  300 samples cross the 256-pointer launch boundary; an odd N and a misaligned source take the
    scalar path; a vector of the wrong length is refused with the reference's message."""
    L = icr.lib()
    rng = np.random.default_rng(12)
    for B, H, W, shift in ((300, 16, 20, 0), (5, 15, 13, 0), (3, 16, 16, 1)):
        N = H * W
        backing = [torch.from_numpy(rng.integers(0, 1 << 30, N + 4, dtype=np.int32)).to(dev) for _ in range(B)]
        maps = [t[shift:shift + N] for t in backing]
        ptrs = (C.c_void_p * B)(*[m.data_ptr() for m in maps])
        sizes = (C.c_int64 * B)(*[N] * B)
        want = torch.stack(maps).view(B, H, W)
        out = torch.zeros((B, H, W), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            rc = L.icr_pack_node_labels(ptrs, sizes, B, C.c_int64(N), _p(out), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, L.icr_last_error()
        assert torch.equal(out, want)
        out2 = torch.zeros_like(out)
        rc = host.icrh_labels_from_node_mappings(ptrs, sizes, B, H, W, 0, _p(out2))
        assert rc == 0, host.icrh_last_error()
        assert torch.equal(out2, want)
    sizes[1] = N - 1
    with torch.cuda.device(dev):
        assert L.icr_pack_node_labels(ptrs, sizes, B, C.c_int64(N), _p(out), None) != 0
    assert "expected N" in L.icr_last_error().decode()
    assert host.icrh_labels_from_node_mappings(ptrs, sizes, B, H, W, 0, _p(out2)) != 0
    assert b"unexpected size" in host.icrh_last_error()
