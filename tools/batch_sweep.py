#!/usr/bin/env python
"""Per-kernel device times of the reward path for batches of 8 .. 4096 config-5 images (inputs resident in HBM, CUDA events
between the library's launches): separates the fixed cost of a launch from the per-image rate.  DESIGN.md section 6."""
import ctypes as C, json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_compression_cpp_b200 import reward, synth
dev = torch.device("cuda:0"); L = reward.lib()
im, lab, sz = synth.make_batch(5, B=64)
if "bg" in sys.argv[1:]:          # a third of each map's rectangles melted into a border-touching background (one big segment per map)
    lab = lab.copy()
    for b in range(lab.shape[0]):
        ids = np.unique(lab[b])
        bg = np.random.default_rng(b).choice(ids, size=len(ids) // 3, replace=False)
        lab[b][np.isin(lab[b], bg)] = -5
        lab[b][0, :], lab[b][-1, :], lab[b][:, 0], lab[b][:, -1] = -5, -5, -5, -5
for B in ((8, 64, 128, 256, 512, 1024) if "bg" in sys.argv[1:] else (8, 64, 128, 256, 512, 1024, 2048, 4096)):
    reps = -(-B // 64)
    ti = torch.from_numpy(np.tile(im, (reps, 1, 1, 1))[:B]).to(dev); tl = torch.from_numpy(np.tile(lab, (reps, 1, 1))[:B]).to(dev); ts = torch.from_numpy(np.tile(sz, reps)[:B]).to(dev)
    ws = reward.Workspace(dev, B, 224, 224, Kmax=1024); prm = reward.make_params(); out = torch.empty(B, dtype=torch.float64, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    def step():
        rc = L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, 224, 224, C.byref(prm), C.c_void_p(out.data_ptr()), ws.handle, st)
        assert rc == 0
    for _ in range(5): step()
    torch.cuda.synchronize()
    L.icr_workspace_set_profile(ws.handle, 1)
    k = np.zeros((20, 4), np.float32)
    for i in range(20):
        step(); L.icr_profile_read(ws.handle, k[i].ctypes.data_as(C.c_void_p))
    L.icr_workspace_set_profile(ws.handle, 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): step()
    e1.record(); torch.cuda.synchronize()
    m = k.mean(0) * 1e3
    print(f"B={B:5d} step {e0.elapsed_time(e1)/20*1e3:8.1f} us  prep {m[0]:7.1f} est {m[1]:7.1f} big {m[2]:5.1f}", flush=True)
    del ws, ti, tl, ts
