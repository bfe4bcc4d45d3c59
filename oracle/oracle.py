"""TEST INFRASTRUCTURE ONLY: ctypes/numpy front end of oracle/liboracle.so (reward_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Params(C.Structure):
    _fields_ = [("min_pixels", C.c_int32), ("L_min", C.c_int32), ("beta", C.c_float),
                ("b_match_token", C.c_float), ("gamma", C.c_float), ("overhead_base", C.c_double),
                ("adaptive_filter", C.c_int32), ("lambda_", C.c_double)]


class Debug(C.Structure):
    _fields_ = [("row_costs", C.c_void_p), ("filter_id", C.c_void_p), ("residuals", C.c_void_p),
                ("hist", C.c_void_p), ("match", C.c_void_p), ("Hbar", C.c_void_p)]


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        src = os.path.join(_HERE, "reward_oracle.c")
        if not os.path.exists(path) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(path)):
            build()
        L = C.CDLL(path)
        L.orc_estimate_segment.restype = C.c_double
        L.orc_compact_labels.restype = C.c_int64
        L.orc_ema_update.restype = C.c_double
        _LIB = L
    return _LIB


def params(min_pixels=1, L_min=4, beta=0.012167, b_match_token=18.0, gamma=0.1,
           overhead_base=9.308622, adaptive_filter=True, lambda_=0.5):
    return Params(min_pixels, L_min, beta, b_match_token, gamma, overhead_base, int(bool(adaptive_filter)), lambda_)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def chw3_f32_to_rgba(chw):
    chw = np.ascontiguousarray(chw, dtype=np.float32)
    _, H, W = chw.shape
    out = np.empty((H, W, 4), np.uint8)
    lib().orc_chw3_f32_to_rgba(_p(chw), C.c_int(H), C.c_int(W), _p(out))
    return out


def compact_labels(labels_hw):
    lab = np.ascontiguousarray(labels_hw, dtype=np.int64)
    comp = np.empty(lab.shape, np.int64)
    uniq = np.empty(lab.size, np.int64)
    K = lib().orc_compact_labels(_p(lab), C.c_int64(lab.size), _p(comp), _p(uniq))
    return comp, uniq[:K].copy()


def segment_stats(compact_hw, K):
    comp = np.ascontiguousarray(compact_hw, dtype=np.int64)
    H, W = comp.shape
    counts = np.empty(K, np.int32)
    bboxes = np.empty((K, 4), np.int32)
    lib().orc_segment_stats(_p(comp), C.c_int(H), C.c_int(W), C.c_int(K), _p(counts), _p(bboxes))
    return counts, bboxes


def slice_segment(rgba, compact_hw, k, x0, y0, w, h):
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    comp = np.ascontiguousarray(compact_hw, dtype=np.int64)
    out = np.empty((h, w, 4), np.uint8)
    lib().orc_slice(_p(rgba), _p(comp), C.c_int(comp.shape[1]), C.c_int64(k),
                    C.c_int(x0), C.c_int(y0), C.c_int(w), C.c_int(h), _p(out))
    return out


def estimate_segment(rgba, compact_hw, k, x0, y0, w, h, prm=None, debug=False):
    """Returns size, or (size, dict of debug tables) when debug=True."""
    prm = prm or params()
    rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
    comp = np.ascontiguousarray(compact_hw, dtype=np.int64)
    H, W = comp.shape
    d = None
    tabs = None
    if debug and w > 0 and h > 0:
        tabs = dict(row_costs=np.zeros((h, 5), np.uint32), filter_id=np.zeros(h, np.uint8),
                    residuals=np.zeros(w * h * 4, np.uint8), hist=np.zeros((4, 256), np.uint32),
                    match=np.zeros(3, np.uint64), Hbar=np.zeros(1, np.float64))
        d = Debug(*[_p(tabs[n]) for n in ("row_costs", "filter_id", "residuals", "hist", "match", "Hbar")])
    s = lib().orc_estimate_segment(_p(rgba), _p(comp), C.c_int(W), C.c_int(H), C.c_int64(k),
                                   C.c_int(x0), C.c_int(y0), C.c_int(w), C.c_int(h),
                                   C.byref(prm), C.byref(d) if d is not None else None)
    return (s, tabs) if debug else s


def compute_rewards(images, labels, image_sizes, prm=None, details=False, Kstride=None):
    """images f32 [B,3,H,W], labels int [B,H,W], image_sizes [B] -> rewards f64 [B]
    (+ dict(K, counts, bboxes, seg_sizes) when details=True)."""
    prm = prm or params()
    images = np.ascontiguousarray(images, dtype=np.float32)
    labels = np.ascontiguousarray(labels, dtype=np.int64)
    sizes = np.ascontiguousarray(image_sizes, dtype=np.float64).reshape(-1)
    B, _, H, W = images.shape
    rewards = np.zeros(B, np.float64)
    if details:
        Ks = int(Kstride or H * W)
        K = np.zeros(B, np.int32)
        counts = np.zeros((B, Ks), np.int32)
        bboxes = np.zeros((B, Ks, 4), np.int32)
        seg = np.zeros((B, Ks), np.float64)
        lib().orc_compute_rewards(_p(images), _p(labels), _p(sizes), C.c_int(B), C.c_int(H), C.c_int(W),
                                  C.byref(prm), _p(rewards), _p(K), C.c_int(Ks), _p(counts), _p(bboxes), _p(seg))
        return rewards, dict(K=K, counts=counts, bboxes=bboxes, seg_sizes=seg)
    lib().orc_compute_rewards(_p(images), _p(labels), _p(sizes), C.c_int(B), C.c_int(H), C.c_int(W),
                              C.byref(prm), _p(rewards), None, C.c_int(0), None, None, None)
    return rewards


def ema_update(value, initialized, momentum, rewards):
    r = np.ascontiguousarray(rewards, dtype=np.float64).reshape(-1)
    init = C.c_int(int(bool(initialized)))
    v = lib().orc_ema_update(C.c_double(value), C.byref(init), C.c_double(momentum), _p(r), C.c_int(r.size))
    return v, bool(init.value)
