#!/usr/bin/env python
"""Times the other BASELINE.json configs (2: 64 x 224^2 rectangles, 3: 256 x 512^2 Voronoi "SLIC-style", 4: 256 x 224^2 with
8 label maps per image) through the C ABI with inputs resident in HBM, and spot-checks each against the CPU oracle.
One JSON line per config.  Not the driver's bench (bench.py is); it shows how the kernels behave off the headline workload."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6552.3


def run(name, images, labels, sizes, pool_reps, steps=10):
    import torch
    from image_compression_cpp_b200 import reward
    from oracle import oracle as orc
    dev = torch.device("cuda:0")
    L = reward.lib()
    multi = labels.ndim == 4
    S = labels.shape[1] if multi else 1
    # tile a pool of distinct images up to the config's batch size
    images_t = np.tile(images, (pool_reps, 1, 1, 1))
    labels_t = np.tile(labels, (pool_reps,) + (1,) * (labels.ndim - 1))
    sizes_t = np.tile(sizes, pool_reps)
    B, _, H, W = images_t.shape
    ti, tl, ts = (torch.from_numpy(a).to(dev) for a in (images_t, labels_t, sizes_t))
    K = [len(np.unique(l)) for l in labels.reshape(-1, H, W)[:32]]
    ws = reward.Workspace(dev, B * S, H, W, Kmax=2048)
    prm = reward.make_params()
    code = 1 if labels.dtype == np.int64 else 0                   # ICR_LABELS_I64 / ICR_LABELS_I32
    out = torch.empty(B * S, dtype=torch.float64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def step():
        rc = L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), code, C.c_void_p(ts.data_ptr()), B, S, H, W,
                                   C.byref(prm), C.c_void_p(out.data_ptr()), ws.handle, st)
        assert rc == 0, L.icr_last_error()
    for _ in range(3):
        step()
    ws.check()
    torch.cuda.synchronize()
    L.icr_workspace_set_profile(ws.handle, 1)
    kms = np.zeros((steps, 4), np.float32)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step()
        L.icr_profile_read(ws.handle, kms[i].ctypes.data_as(C.c_void_p))
    e1.record()
    torch.cuda.synchronize()
    L.icr_workspace_set_profile(ws.handle, 0)
    k = kms.mean(0)
    ms = float(k.sum())
    # spot parity on the first maps
    n = min(4, images.shape[0])
    if multi:
        want = orc.compute_rewards(np.repeat(images[:n], S, axis=0), labels[:n].reshape(n * S, H, W), np.repeat(sizes[:n], S))
    else:
        want = orc.compute_rewards(images[:n], labels[:n], sizes[:n])
    got = out.cpu().numpy()[:want.size]
    err = float(np.max(np.abs(got - want) / np.abs(want)))
    px_bytes = 12 + labels.dtype.itemsize * S
    alg = B * (H * W * px_bytes + S * (np.mean(K) * 28 + 16))
    print(json.dumps({"config": name, "images": B, "label_maps": B * S, "H": H, "W": W, "mean_K": float(np.mean(K)),
                      "ms_per_step_kernels": ms, "images_per_sec": B / (ms * 1e-3), "rewards_per_sec": B * S / (ms * 1e-3),
                      "kernel_ms": {"prep": float(k[0]), "narrow": float(k[1]), "big": float(k[2])},
                      "algorithmic_GBps": alg / (ms * 1e-3) / 1e9, "max_rel_err_vs_oracle": err}))
    assert err < 1e-5, err
    # the row after the path: every slice of the batch into one packed buffer (plan + emit), tables already in place
    offsets = torch.zeros(B * S * 2048 + 1, dtype=torch.int64, device=dev)
    summary = torch.zeros(2, dtype=torch.int64, device=dev)
    L.icr_slices_plan(ws.handle, C.c_void_p(offsets.data_ptr()), C.c_longlong(offsets.numel()), C.c_void_p(summary.data_ptr()), st)
    nseg, nbytes = (int(x) for x in summary.tolist())
    buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    tp = te = 0.0
    for i in range(steps + 2):
        ev[0].record()
        L.icr_slices_plan(ws.handle, C.c_void_p(offsets.data_ptr()), C.c_longlong(offsets.numel()), C.c_void_p(summary.data_ptr()), st)
        ev[1].record()
        rc = L.icr_slices_emit(ws.handle, C.c_void_p(offsets.data_ptr()), C.c_void_p(buf.data_ptr()), C.c_size_t(nbytes), 0, st)
        ev[2].record()
        assert rc == 0, L.icr_last_error()
        torch.cuda.synchronize()
        if i >= 2:
            tp += ev[0].elapsed_time(ev[1]) / steps
            te += ev[1].elapsed_time(ev[2]) / steps
    # HBM bytes the emit has to move: every input plane once (images are shared by the S maps of an image) + the slices written.
    # "logical" counts 16 B read per bbox pixel instead: overlapping bboxes re-read pixels, which the L2 should serve.
    hbm_bytes = B * H * W * 12 + B * S * H * W * tl.element_size() + nbytes
    print(json.dumps({"config": name + " -- slices", "segments": nseg, "slice_bytes": nbytes, "plan_ms": tp, "emit_ms": te,
                      "emit_hbm_GBps": hbm_bytes / (te * 1e-3) / 1e9, "emit_frac_of_hbm_peak": hbm_bytes / (te * 1e-3) / 1e9 / hbm_peak(),
                      "emit_logical_GBps": (nbytes * 5) / (te * 1e-3) / 1e9, "images_per_sec": B / ((tp + te) * 1e-3),
                      "note": "hbm bytes = every input plane once + 4 B written per bbox pixel; logical = 16 B read + 4 B written per bbox pixel"}))
    # the decoder row: every slice of a label map pasted back into its canvas (one clear + one launch per map)
    nmaps = min(B * S, 64)
    canvas = torch.empty((nmaps, H, W, 4), dtype=torch.uint8, device=dev)
    td = 0.0
    for i in range(steps + 2):
        ev[0].record()
        for v in range(nmaps):
            rc = L.icr_reassemble_map(ws.handle, v, C.c_void_p(offsets.data_ptr()), C.c_void_p(buf.data_ptr()), C.c_void_p(canvas[v].data_ptr()), st)
        ev[1].record()
        assert rc == 0, L.icr_last_error()
        torch.cuda.synchronize()
        if i >= 2:
            td += ev[0].elapsed_time(ev[1]) / steps
    # ... and the whole batch in one clear + one launch
    all_canvas = torch.empty((B * S, H, W, 4), dtype=torch.uint8, device=dev)
    tb = 0.0
    for i in range(steps + 2):
        ev[0].record()
        rc = L.icr_reassemble_batch(ws.handle, C.c_void_p(offsets.data_ptr()), C.c_void_p(buf.data_ptr()), C.c_void_p(all_canvas.data_ptr()), st)
        ev[1].record()
        assert rc == 0, L.icr_last_error()
        torch.cuda.synchronize()
        if i >= 2:
            tb += ev[0].elapsed_time(ev[1]) / steps
    assert torch.equal(all_canvas[:nmaps], canvas)
    dec_bytes = nbytes + 2 * all_canvas.numel()                   # slices read, canvases cleared and written
    want_img = orc.chw3_f32_to_rgba(images[0])
    ok = bool(np.array_equal(canvas[0].cpu().numpy(), want_img))
    print(json.dumps({"config": name + " -- decoder", "maps": nmaps, "ms_per_map": td / nmaps, "maps_per_sec": nmaps / (td * 1e-3),
                      "canvas_GBps": nmaps * H * W * 4 / (td * 1e-3) / 1e9,
                      "batch_ms": tb, "batch_maps_per_sec": B * S / (tb * 1e-3), "batch_hbm_GBps": dec_bytes / (tb * 1e-3) / 1e9,
                      "batch_frac_of_hbm_peak": dec_bytes / (tb * 1e-3) / 1e9 / hbm_peak(), "round_trip_of_map_0_matches_rule_C_image": ok,
                      "note": "icr_reassemble_map per map: clear + paste of all its slices; icr_reassemble_batch: all maps, hbm bytes = slices read + canvases cleared and written"}))
    assert ok
    ws.close()


def main():
    from image_compression_cpp_b200 import synth
    which = sys.argv[1:] or ["2", "3", "4"]
    if "2" in which:
        im, lab, sz = synth.make_batch(2)
        run("cfg2: 64 x 224x224, BSP rectangles", im, lab, sz, 1)
    if "3" in which:
        im, lab, sz = synth.make_batch(3, B=32)
        run("cfg3: 256 x 512x512, Voronoi (SLIC-style) ~1024 segments", im, lab, sz, 8)
    if "i64" in which:          # the bench workload (512 of config 5) with int64 labels: what compute_rewards_batched's own callers hold
        im, lab, sz = synth.make_batch(5, B=64)
        run("i64: 512 x 224x224, BSP rectangles, int64 labels", im, lab.astype(np.int64), sz, 8)
    if "big" in which:          # adversarial for the warp path: every map is ONE whole-image segment (Rule M, chunk = 4), all through big_kernel
        im, lab, sz = synth.make_batch(2)
        run("big: 512 x 224x224, one whole-image segment per map", im, np.zeros_like(lab), sz, 8)
    if "bg" in which:           # what multicut output looks like: compact segments on a background that touches all four borders
        im, lab, sz = synth.make_batch(5, B=64)
        lab = lab.copy()
        for b in range(lab.shape[0]):
            ids = np.unique(lab[b])
            bg = np.random.default_rng(b).choice(ids, size=len(ids) // 3, replace=False)
            lab[b][np.isin(lab[b], bg)] = -5
            lab[b][0, :], lab[b][-1, :], lab[b][:, 0], lab[b][:, -1] = -5, -5, -5, -5
        run("bg: 512 x 224x224, BSP rectangles on a border-touching background segment (a third of the rectangles)", im, lab, sz, 8)
    if "4" in which:
        im, lab, sz = synth.make_multi_batch(B=32, S=8)
        run("cfg4: 256 x 224x224, 8 label maps per image (rectangles / region growing)", im, lab, sz, 8)


if __name__ == "__main__":
    main()
