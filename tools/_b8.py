import ctypes as C, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_compression_cpp_b200 import reward, synth
dev = torch.device("cuda:0"); L = reward.lib()
B = 8
im, lab, sz = synth.make_batch(5, B=B)
ti, tl, ts = (torch.from_numpy(a).to(dev) for a in (im, lab, sz))
ws = reward.Workspace(dev, B, 224, 224, Kmax=1024); prm = reward.make_params(); out = torch.empty(B, dtype=torch.float64, device=dev)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(6):
    rc = L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, 224, 224, C.byref(prm), C.c_void_p(out.data_ptr()), ws.handle, st)
    assert rc == 0
torch.cuda.synchronize()
print("ok")
