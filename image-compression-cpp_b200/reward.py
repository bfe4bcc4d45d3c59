"""Host-side mirror of the reference's call surface for the reward path, over the C ABI.

Same names, argument meaning and error behaviour as the reference's three public functions
(include/compute_rewards.cuh:5-17, include/segment_stats.cuh:6-10,
include/png_size_estimator.cuh:61-71), with torch tensors used only as device-memory handles:
every call goes through libicr_b200.so (include/icr_b200.h) with raw pointers and the current
CUDA stream.  There is no CPU fallback: a missing library or a CPU tensor where the reference
demands CUDA raises.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ICR_LIB") or os.path.join(_HERE, "libicr_b200.so")       # ICR_LIB: an experimental build (tools/)
_LIB = None

ICR_LABELS_I32, ICR_LABELS_I64 = 0, 1


class IcrParams(C.Structure):
    _fields_ = [("min_pixels", C.c_int32), ("L_min", C.c_int32), ("beta", C.c_float),
                ("b_match_token", C.c_float), ("gamma", C.c_float), ("overhead_base", C.c_double),
                ("adaptive_filter", C.c_int32), ("lambda_", C.c_double)]


SYMBOLS = ["icr_last_error", "icr_version", "icr_default_params", "icr_workspace_bytes", "icr_workspace_create",
           "icr_workspace_destroy", "icr_compute_rewards", "icr_compute_rewards_gather", "icr_compute_rewards_host", "icr_compute_rewards_host_u8", "icr_check", "icr_check_detail",
           "icr_label_capacity", "icr_segment_stats", "icr_counts_bboxes_from_compact_labels",
           "icr_estimate_png_size_segment", "icr_estimate_png_size_segment_noalloc", "icr_workspace_tables", "icr_debug_segment", "icr_slice_segment",
           "icr_slices_prepare", "icr_slices_plan", "icr_slices_emit", "icr_reassemble_map", "icr_reassemble_batch", "icr_paste_slice", "icr_ema_update", "icr_ema_update_gathered", "icr_reward_sum_count", "icr_selftest", "icr_launch_count", "icr_workspace_set_profile",
           "icr_profile_read", "icr_pack_node_labels", "icr_nccl_unique_id", "icr_nccl_comm_create", "icr_nccl_comm_destroy", "icr_nccl_comm_ranks", "icr_exchange_rewards"]


def lib():
    """The C-ABI library; raises if it has not been built (python -c 'import __graft_entry__ as g; g.build()')."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with __graft_entry__.build(); there is no fallback path")
        L = C.CDLL(LIB_PATH)
        L.icr_last_error.restype = C.c_char_p
        L.icr_workspace_bytes.restype = C.c_size_t
        L.icr_label_capacity.restype = C.c_int64
        L.icr_launch_count.restype = C.c_longlong
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise RuntimeError(lib().icr_last_error().decode())


def make_params(min_pixels_per_segment=1, L_min=4, beta=0.012167, b_match_token=18.0, gamma=0.1,
                overhead_base=9.308622, adaptive_filter=True, lambda_=0.5):
    return IcrParams(int(min_pixels_per_segment), int(L_min), float(beta), float(b_match_token), float(gamma),
                     float(overhead_base), int(bool(adaptive_filter)), float(lambda_))


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _d2d(dst, src_ptr):
    """Device-to-device copy from a raw workspace pointer into a torch tensor, on the current stream
    (the CUDA runtime torch already loaded)."""
    rt = C.CDLL("libcudart.so.12")
    rc = rt.cudaMemcpyAsync(_ptr(dst), src_ptr, C.c_size_t(dst.numel() * dst.element_size()), C.c_int(3), _stream())
    if rc != 0:
        raise RuntimeError(f"cudaMemcpyAsync failed with {rc}")


class Workspace:
    """Device scratch for V label maps of HxW pixels with at most Kmax segments each."""

    def __init__(self, device, V, H, W, Kmax=None):
        self.device = torch.device(device)
        self.V, self.H, self.W = int(V), int(H), int(W)
        self.Kmax = int(min(Kmax or H * W, H * W))
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _check(lib().icr_workspace_create(C.byref(self._h), C.c_int(self.device.index or 0), C.c_int(self.V),
                                              C.c_int(self.H), C.c_int(self.W), C.c_int(self.Kmax)))

    @property
    def handle(self):
        return self._h

    def nbytes(self):
        return lib().icr_workspace_bytes(C.c_int(self.V), C.c_int(self.H), C.c_int(self.W), C.c_int(self.Kmax))

    def check(self):
        """One stream sync: raises if a label map of the calls so far could not be processed."""
        bad = C.c_int(-1)
        with torch.cuda.device(self.device):
            _check(lib().icr_check(self._h, _stream(), C.byref(bad)))

    def tables(self):
        """(K [V], counts [V,Kmax], bboxes [V,Kmax,4], seg_sizes [V,Kmax]) copied into fresh torch tensors."""
        pK, pc, pb, ps, pp = (C.c_void_p() for _ in range(5))
        km = C.c_int()
        _check(lib().icr_workspace_tables(self._h, C.byref(pK), C.byref(pc), C.byref(pb), C.byref(ps), C.byref(pp), C.byref(km)))
        V, Kmax = self.V, self.Kmax
        out = []
        for p, shape, dt in ((pK, (V,), torch.int32), (pc, (V, Kmax), torch.int32), (pb, (V, Kmax, 4), torch.int32),
                             (ps, (V, Kmax), torch.float64)):
            t = torch.empty(shape, dtype=dt, device=self.device)
            _d2d(t, p)
            out.append(t)
        return tuple(out)

    def keys(self):
        """[V,Kmax] int64: original label value of each compact id (ascending within a map: Rule L)."""
        pk = C.c_void_p()
        _check(lib().icr_workspace_tables(self._h, None, None, None, None, C.byref(pk), None))
        t = torch.empty((self.V, self.Kmax), dtype=torch.int64, device=self.device)
        _d2d(t, pk)
        return t

    def debug_segment(self, v, k, h, params=None):
        """Row costs [h,5], filter ids [h], histogram [4,256], match [3], size of segment k of map v."""
        params = params or make_params()
        dev = self.device
        rc = torch.zeros((h, 5), dtype=torch.int32, device=dev)
        fi = torch.zeros(h, dtype=torch.uint8, device=dev)
        hist = torch.zeros((4, 256), dtype=torch.int32, device=dev)
        match = torch.zeros(3, dtype=torch.int64, device=dev)
        size = torch.zeros(1, dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            _check(lib().icr_debug_segment(self._h, C.c_int(v), C.c_int(k), C.byref(params), _ptr(rc), _ptr(fi), _ptr(hist),
                                           _ptr(match), _ptr(size), _stream()))
        return rc, fi, hist, match, size

    def slice_segment(self, v, k, w, h):
        out = torch.zeros((h, w, 4), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().icr_slice_segment(self._h, C.c_int(v), C.c_int(k), _ptr(out), C.c_size_t(out.numel()), _stream()))
        return out

    def slices(self, order=0):
        """All slices of the last batch call in one packed device buffer (src/image_slicer.cpp:81-125 for the whole batch).
        Returns (buffer u8 [total bytes], offsets int64 [segments + 1]); slice f = (segments of earlier maps) + k is
        buffer[offsets[f]:offsets[f+1]] viewed as [h,w,4].  One 16-byte read-back sizes the buffer."""
        dev = self.device
        offsets = torch.zeros(self.V * self.Kmax + 1, dtype=torch.int64, device=dev)
        summary = torch.zeros(2, dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            _check(lib().icr_slices_plan(self._h, _ptr(offsets), C.c_longlong(offsets.numel()), _ptr(summary), _stream()))
            nseg, nbytes = (int(x) for x in summary.tolist())
            buf = torch.empty(max(nbytes, 4), dtype=torch.uint8, device=dev)
            _check(lib().icr_slices_emit(self._h, _ptr(offsets), _ptr(buf), C.c_size_t(nbytes), C.c_int(order), _stream()))
        return buf[:nbytes], offsets[:nseg + 1]

    def reassemble_map(self, v, buf, offsets):
        """Decoder check (src/reassemble.cpp:50-107): canvas u8 [H,W,4] with every slice of map v pasted where alpha > 0."""
        canvas = torch.empty((self.H, self.W, 4), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().icr_reassemble_map(self._h, C.c_int(v), _ptr(offsets), _ptr(buf), _ptr(canvas), _stream()))
        return canvas

    def reassemble_batch(self, buf, offsets):
        """All V label maps of the last batch at once: canvases u8 [V,H,W,4], one clear and one launch."""
        canvases = torch.empty((self.V, self.H, self.W, 4), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _check(lib().icr_reassemble_batch(self._h, _ptr(offsets), _ptr(buf), _ptr(canvases), _stream()))
        return canvases

    def close(self):
        if self._h:
            lib().icr_workspace_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_WS_CACHE = {}


def get_workspace(device, V, H, W, Kmax=None):
    """Grow-only cache, one workspace per (device, H, W): the role of get_png_ws (src/png_size_estimator.cu:466-469)."""
    device = torch.device(device)
    key = (device.index or 0, H, W)
    Kmax = int(min(Kmax or H * W, H * W))
    ws = _WS_CACHE.get(key)
    if ws is None or ws.V < V or ws.Kmax < Kmax:
        if ws is not None:
            torch.cuda.synchronize(device)
            ws.close()
        ws = Workspace(device, max(V, ws.V if ws else 0), H, W, max(Kmax, ws.Kmax if ws else 0))
        _WS_CACHE[key] = ws
    return ws


def compute_rewards_batched(images_bchw_f32, labels, image_sizes, min_pixels_per_segment=1, L_min=4,
                            beta=0.012167, b_match_token=18.0, gamma=0.1, overhead_base=9.308622,
                            adaptive_filter=True, lambda_=0.5, workspace=None, check=True):
    """Drop-in for compute_rewards_batched (src/compute_rewards.cu:55-196): rewards f64 [B] on labels' device.

    labels may also be [B,S,H,W] (S label maps per image, one image read): rewards are then [B,S]."""
    if images_bchw_f32.dim() != 4 or images_bchw_f32.size(1) != 3:
        raise RuntimeError("images must be [B,3,H,W]")
    if images_bchw_f32.dtype != torch.float32:
        raise RuntimeError("images must be float32 (CV_32FC3)")
    if not labels.is_cuda:
        raise RuntimeError("labels must be CUDA")
    if not image_sizes.is_cuda:
        raise RuntimeError("image_sizes must be CUDA")
    if labels.dim() not in (3, 4):
        raise RuntimeError("labels must be [B,H,W]")
    B, _, H, W = images_bchw_f32.shape
    if image_sizes.numel() != B:
        raise RuntimeError("image_sizes must have B elements")
    multi = labels.dim() == 4
    S = labels.size(1) if multi else 1
    if labels.size(0) != B or labels.size(-2) != H or labels.size(-1) != W:
        raise RuntimeError("labels must match images: [B,H,W]")
    if labels.dtype.is_floating_point or labels.dtype == torch.bool:
        raise RuntimeError("labels must be an integer tensor")
    dev = labels.device
    with torch.no_grad(), torch.cuda.device(dev):
        images = images_bchw_f32.to(dev, non_blocking=True).contiguous()
        if labels.dtype not in (torch.int32, torch.int64):
            labels = labels.to(torch.int64)
        labels = labels.contiguous()
        sizes = image_sizes.reshape(-1).to(torch.float64).contiguous()
        rewards = torch.empty(B * S, dtype=torch.float64, device=dev)
        ws = workspace or get_workspace(dev, B * S, H, W)
        prm = make_params(min_pixels_per_segment, L_min, beta, b_match_token, gamma, overhead_base, adaptive_filter, lambda_)
        _check(lib().icr_compute_rewards(_ptr(images), _ptr(labels),
                                         C.c_int(ICR_LABELS_I64 if labels.dtype == torch.int64 else ICR_LABELS_I32),
                                         _ptr(sizes), C.c_int(B), C.c_int(S), C.c_int(H), C.c_int(W), C.byref(prm),
                                         _ptr(rewards), ws.handle, _stream()))
        if check:
            ws.check()
    return rewards.view(B, S) if multi else rewards


def compute_counts_bboxes_from_compact_labels_cuda(labels_compact_hw_i64, counts_k_i32, bboxes_k4_i32):
    """Drop-in for src/segment_stats.cu:44-77 (same checks, same in-place outputs)."""
    if not labels_compact_hw_i64.is_cuda:
        raise RuntimeError("labels must be CUDA")
    if labels_compact_hw_i64.dtype != torch.int64:
        raise RuntimeError("labels must be int64")
    if labels_compact_hw_i64.dim() != 2:
        raise RuntimeError("labels must be [H,W]")
    if not (counts_k_i32.is_cuda and bboxes_k4_i32.is_cuda):
        raise RuntimeError("outputs must be CUDA")
    if counts_k_i32.dtype != torch.int32:
        raise RuntimeError("counts must be int32")
    if bboxes_k4_i32.dtype != torch.int32:
        raise RuntimeError("bboxes must be int32")
    if bboxes_k4_i32.dim() != 2 or bboxes_k4_i32.size(1) != 4:
        raise RuntimeError("bboxes must be [K,4]")
    lab = labels_compact_hw_i64.contiguous()
    H, W = lab.shape
    with torch.cuda.device(lab.device):
        _check(lib().icr_counts_bboxes_from_compact_labels(_ptr(lab), C.c_int(H), C.c_int(W), C.c_int(counts_k_i32.numel()),
                                                           _ptr(counts_k_i32), _ptr(bboxes_k4_i32), _stream()))


def estimate_png_size_masked_segment_to_output(img_hwc_u8, labels_compact_hw, counts_k, seg_id_k, x0, y0, w, h,
                                               min_pixels=1, L_min=4, beta=0.012167, b_match_token=18.0, gamma=0.1,
                                               overhead_base=9.308622, adaptive_filter=True, out=None, debug=False):
    """Drop-in for src/png_size_estimator.cu:472-591 on tensors: writes one f64 into `out` (device, 1 element).

    With debug=True also returns (row_costs [h,5], filter_id [h], hist [C,256], match [3])."""
    if img_hwc_u8 is None or labels_compact_hw is None or counts_k is None:
        raise RuntimeError("null ptr")
    full_H, full_W, channels = img_hwc_u8.shape
    dev = img_hwc_u8.device
    out = out if out is not None else torch.zeros(1, dtype=torch.float64, device=dev)
    tabs = None
    ptrs = [None, None, None, None]
    if debug and w > 0 and h > 0:
        tabs = (torch.zeros((h, 5), dtype=torch.int32, device=dev), torch.zeros(h, dtype=torch.uint8, device=dev),
                torch.zeros((channels, 256), dtype=torch.int32, device=dev), torch.zeros(3, dtype=torch.int64, device=dev))
        ptrs = [_ptr(t) for t in tabs]
    with torch.cuda.device(dev):
        _check(lib().icr_estimate_png_size_segment(
            _ptr(img_hwc_u8.contiguous()), C.c_int(full_W), C.c_int(full_H), C.c_int(channels),
            _ptr(labels_compact_hw.contiguous()), _ptr(counts_k), C.c_int64(seg_id_k), C.c_int(x0), C.c_int(y0), C.c_int(w),
            C.c_int(h), C.c_int32(min_pixels), C.c_int(L_min), C.c_float(beta), C.c_float(b_match_token), C.c_float(gamma),
            C.c_double(overhead_base), C.c_int(int(bool(adaptive_filter))), _ptr(out), *ptrs, _stream()))
    return (out, tabs) if debug else out


def segment_stats(labels, Kmax=None, workspace=None, compact=False):
    """Batched Rule L + Rule S: labels [V,H,W] int32/int64 CUDA -> (K [V], counts [V,Kmax], bboxes [V,Kmax,4])
    (+ compact ids int32 [V,H,W], the inverse at::_unique would return, when compact=True)."""
    if not labels.is_cuda:
        raise RuntimeError("labels must be CUDA")
    V, H, W = labels.shape
    if labels.dtype not in (torch.int32, torch.int64):
        labels = labels.to(torch.int64)
    labels = labels.contiguous()
    ws = workspace or get_workspace(labels.device, V, H, W, Kmax)
    comp = torch.empty((V, H, W), dtype=torch.int32, device=labels.device) if compact else None
    with torch.cuda.device(labels.device):
        _check(lib().icr_segment_stats(_ptr(labels), C.c_int(ICR_LABELS_I64 if labels.dtype == torch.int64 else ICR_LABELS_I32),
                                       C.c_int(V), C.c_int(H), C.c_int(W), None, None, None,
                                       _ptr(comp) if compact else None, ws.handle, _stream()))
        ws.check()
    K, counts, bboxes, _ = ws.tables()
    return (K[:V], counts[:V], bboxes[:V], comp) if compact else (K[:V], counts[:V], bboxes[:V])


def labels_from_rama(node_labels_bn_i32, H, W):
    """The label-map handoff of src/fcn/training.cpp:169-172: rama_torch_batched returns int32 [B,N=H*W] CUDA
    contiguous node ids (src/rama_wrapper.cu:99); the reward path takes it as [B,H,W] without a copy or a cast."""
    t = node_labels_bn_i32
    if not t.is_cuda:
        raise RuntimeError("node labels must be a CUDA tensor")
    if t.dtype != torch.int32:
        raise RuntimeError("node labels must be int32")
    if t.dim() != 2 or t.size(1) != H * W:
        raise RuntimeError("node labels must be [B, H*W]")
    if not t.is_contiguous():
        raise RuntimeError("node labels must be contiguous")
    return t.view(t.size(0), H, W)


class EMABaseline:
    """include/ema_baseline.hpp:4-22 on the device: value = m*value + (1-m)*mean(rewards), first call copies."""

    def __init__(self, momentum=0.99):
        self.momentum = float(momentum)
        self.state = None       # device f64 [2] = {value, initialized}

    @property
    def initialized(self):
        return self.state is not None and bool(self.state[1].item() != 0.0)

    @property
    def value(self):
        return self.state[0] if self.state is not None else None

    def update(self, rewards, sum_count=None, advantages=False):
        """rewards f64 [n] CUDA.  sum_count: optional device f64 [2] = global {sum, count} (after an all-reduce).
        Returns value (0-dim view), or (value, advantages [n]) with the standardisation of src/fcn/training.cpp:178-181."""
        if not rewards.is_cuda or rewards.dtype != torch.float64:
            raise RuntimeError("rewards must be a CUDA float64 tensor")
        r = rewards.reshape(-1).contiguous()
        if self.state is None:
            self.state = torch.zeros(2, dtype=torch.float64, device=r.device)
        adv = torch.empty_like(r) if advantages else None
        with torch.cuda.device(r.device):
            _check(lib().icr_ema_update(_ptr(r), C.c_int(r.numel()), _ptr(sum_count) if sum_count is not None else None,
                                        C.c_double(self.momentum), _ptr(self.state), _ptr(adv) if advantages else None, _stream()))
        return (self.state[0], adv.view_as(rewards)) if advantages else self.state[0]


def paste_slice(canvas, slice_hw4, x, y):
    """canvas u8 [H,W,4] (CUDA) receives slice u8 [h,w,4] at (x, y) where the slice's alpha > 0, clipped to the canvas."""
    H, W = canvas.shape[:2]
    h, w = slice_hw4.shape[:2]
    with torch.cuda.device(canvas.device):
        _check(lib().icr_paste_slice(_ptr(canvas), C.c_int(W), C.c_int(H), _ptr(slice_hw4.contiguous()), C.c_int(x), C.c_int(y),
                                     C.c_int(w), C.c_int(h), _stream()))
    return canvas


def ema_update_gathered(ema, peer, advantages=False):
    """Rule E on the gathered vector of sharding.PeerGather's current step: the kernel first waits for every rank's flag."""
    r = peer.gathered()
    if ema.state is None:
        ema.state = torch.zeros(2, dtype=torch.float64, device=r.device)
    adv = torch.empty_like(r) if advantages else None
    with torch.cuda.device(r.device):
        _check(lib().icr_ema_update_gathered(_ptr(r), C.c_int(r.numel()), _ptr(peer.flags), C.c_int(peer.world),
                                             C.c_ulonglong(peer.step), C.c_double(ema.momentum), _ptr(ema.state),
                                             _ptr(adv) if advantages else None, _stream()))
    return (ema.state[0], adv) if advantages else ema.state[0]


def reward_sum_count(rewards):
    r = rewards.reshape(-1).contiguous()
    out = torch.empty(2, dtype=torch.float64, device=r.device)
    with torch.cuda.device(r.device):
        _check(lib().icr_reward_sum_count(_ptr(r), C.c_int(r.numel()), _ptr(out), _stream()))
    return out


def selftest():
    n = C.c_int(-1)
    _check(lib().icr_selftest(C.byref(n)))
    return n.value
