"""Generates tests/golden/reference_cases.npz ON THE GPU BOX from the reference's own CUDA sources
(oracle/_ref/libref_reward.so = /root/reference/src/{segment_stats,png_size_estimator,compute_rewards}.cu
compiled unchanged by oracle/build_ref.sh).  Run:  gpurun -- 'python tests/golden/make_golden.py gpurun_out/reference_cases.npz'
then copy the file to tests/golden/.  Inputs are stored with the outputs so the fixture is self-contained."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from image_compression_cpp_b200 import synth  # noqa: E402


def ref_lib():
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_reward.so"))
    L.ref_last_error.restype = C.c_char_p
    return L


def ref_rewards(L, images, labels, sizes, min_pixels=1, L_min=4, beta=0.012167, token=18.0, gamma=0.1, overhead=9.308622,
                adaptive=1, lam=0.5):
    dev = torch.device("cuda:0")
    ti, ts = torch.from_numpy(images).to(dev), torch.from_numpy(sizes).to(dev)
    tl = torch.from_numpy(labels).to(dev)
    B, _, H, W = images.shape
    out = torch.empty(B, dtype=torch.float64, device=dev)
    rc = L.ref_compute_rewards(C.c_void_p(ti.data_ptr()), 1, C.c_void_p(tl.data_ptr()), int(labels.dtype == np.int64),
                               C.c_void_p(ts.data_ptr()), B, H, W, 0, min_pixels, L_min, C.c_float(beta), C.c_float(token),
                               C.c_float(gamma), C.c_double(overhead), adaptive, C.c_double(lam), C.c_void_p(out.data_ptr()))
    if rc:
        raise RuntimeError(L.ref_last_error().decode())
    return out.cpu().numpy()


def ref_stats_and_sizes(L, rgba, compact, K, prm):
    """counts, bboxes and per-segment sizes of one image from the reference's two inner entry points."""
    dev = torch.device("cuda:0")
    H, W = compact.shape
    tl = torch.from_numpy(compact.astype(np.int64)).to(dev)
    timg = torch.from_numpy(rgba).to(dev)
    cnt = torch.empty(K, dtype=torch.int32, device=dev)
    bb = torch.empty((K, 4), dtype=torch.int32, device=dev)
    if L.ref_segment_stats(C.c_void_p(tl.data_ptr()), H, W, K, 0, C.c_void_p(cnt.data_ptr()), C.c_void_p(bb.data_ptr())):
        raise RuntimeError(L.ref_last_error().decode())
    bbh = bb.cpu().numpy()
    sz = torch.zeros(K, dtype=torch.float64, device=dev)
    for k in range(K):
        x0, y0, x1, y1 = [int(v) for v in bbh[k]]
        if L.ref_estimate_segment(C.c_void_p(timg.data_ptr()), W, H, 4, C.c_void_p(tl.data_ptr()), C.c_void_p(cnt.data_ptr()),
                                  C.c_int64(k), x0, y0, x1 - x0 + 1, y1 - y0 + 1, 0, prm["min_pixels"], prm["L_min"],
                                  C.c_float(prm["beta"]), C.c_float(prm["token"]), C.c_float(prm["gamma"]),
                                  C.c_double(prm["overhead"]), prm["adaptive"], C.c_void_p(sz.data_ptr() + 8 * k)):
            raise RuntimeError(L.ref_last_error().decode())
    return cnt.cpu().numpy(), bbh, sz.cpu().numpy()


def cases():
    """name -> (images f32 [B,3,H,W], labels, sizes, params)"""
    d = dict(min_pixels=1, L_min=4, beta=0.012167, token=18.0, gamma=0.1, overhead=9.308622, adaptive=1, lam=0.5)
    out = {}
    im, lab, sz = synth.make_batch(2, B=6)
    out["cfg2_rect_224"] = (im, lab, sz, dict(d))
    H = W = 96
    im = np.stack([synth.image_f32_chw(b, H, W, 50 + b) for b in range(6)])
    lab = np.stack([synth.rect_partition(H, W, split_prob=0.9, min_rect_count=20, seed=60 + b) * 5 - 17 for b in range(6)]).astype(np.int32)
    out["fine_rect_96"] = (im, lab, np.full(6, 9000.0), dict(d))
    H, W = 72, 88
    im = np.stack([synth.image_f32_chw(b + 2, H, W, 70 + b) for b in range(4)])
    lab = np.stack([synth.bfs_partition(H, W, 11, seed=1), synth.voronoi_partition(H, W, 16, 6, seed=2),
                    synth.bfs_partition(H, W, 3, seed=3), synth.voronoi_partition(H, W, 9, 3, seed=4)]).astype(np.int64)
    out["bfs_voronoi_i64"] = (im, lab, np.full(4, 20000.0), dict(d))
    out["non_adaptive"] = (im, lab, np.full(4, 20000.0), dict(d, adaptive=0))
    out["min_pixels_50_lmin_2"] = (im, lab, np.full(4, 20000.0), dict(d, min_pixels=50, L_min=2, lam=0.25))
    H = W = 224                         # whole-image bbox: chunk = 4 bytes, Rule M is live
    im = np.stack([synth.image_f32_chw(f, H, W, 8 + f) for f in (1, 2, 3)])
    lab = np.zeros((3, H, W), np.int32)
    lab[:, 60:90, 100:140] = 1
    lab[2] = synth.bfs_partition(H, W, 3, seed=2)
    out["rule_m_224"] = (im, lab, np.full(3, 70000.0), dict(d))
    out["rule_m_224_lmin1"] = (im[:2], lab[:2], np.full(2, 70000.0), dict(d, L_min=1))
    rng = np.random.default_rng(3)
    im = rng.random((2, 3, 64, 64), dtype=np.float32)
    lab = np.stack([synth.grid_partition(64, 64, 16, 16)] * 2).astype(np.int32)
    out["unit_range_floats"] = (im, lab, np.full(2, 5000.0), dict(d))
    base = rng.integers(-20, 280, (2, 3, 32, 32)).astype(np.float32) + np.float32(0.5)
    out["half_rounding_clamp"] = (base, np.stack([synth.grid_partition(32, 32, 8, 16)] * 2).astype(np.int32), np.full(2, 2000.0), dict(d))
    out["single_segment"] = (np.stack([synth.image_f32_chw(2, 48, 48, 5)]), np.full((1, 48, 48), 9, np.int32), np.array([4000.0]), dict(d))
    return out


def main(path):
    sys.path.insert(0, ROOT)
    from oracle import oracle as orc
    L = ref_lib()
    blob = {}
    for name, (im, lab, sz, prm) in cases().items():
        r = ref_rewards(L, im, lab, sz, prm["min_pixels"], prm["L_min"], prm["beta"], prm["token"], prm["gamma"], prm["overhead"],
                        prm["adaptive"], prm["lam"])
        blob[f"{name}/images"] = im
        blob[f"{name}/labels"] = lab
        blob[f"{name}/sizes"] = sz
        blob[f"{name}/params"] = np.array([prm[k] for k in ("min_pixels", "L_min", "beta", "token", "gamma", "overhead", "adaptive", "lam")], np.float64)
        blob[f"{name}/rewards"] = r
        # inner entry points on image 0 (compact ids from the oracle's Rule L == at::_unique's inverse)
        rgba = orc.chw3_f32_to_rgba(im[0])
        comp, uniq = orc.compact_labels(lab[0])
        cnt, bb, ssz = ref_stats_and_sizes(L, rgba, comp, len(uniq), prm)
        blob[f"{name}/counts0"], blob[f"{name}/bboxes0"], blob[f"{name}/seg_sizes0"] = cnt, bb, ssz
        print(name, r[:3], len(uniq))
    np.savez_compressed(path, **blob)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "reference_cases.npz"))
