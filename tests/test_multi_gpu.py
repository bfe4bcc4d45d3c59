"""Two GPUs of one box (skipped on a single-GPU box): the fused exchange -- rewards stored by the reward kernel straight into
every rank's gathered vector through peer memory -- against the plain NCCL all-gather, and the EMA on the gathered vector
against the single-process value."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import ctypes as C, os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch, torch.distributed as dist
from image_compression_cpp_b200 import reward, sharding, synth
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B = 6                                             # per rank
images, labels, sizes = synth.make_batch(2, B=B * world)
lo, hi = sharding.shard_range(B * world, world, rank)
ti, tl, ts = (torch.from_numpy(np.ascontiguousarray(a[lo:hi])).to(dev) for a in (images, labels, sizes))
L = reward.lib(); prm = reward.make_params(); ws = reward.Workspace(dev, B, 224, 224, None)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
peer = sharding.PeerGather(B, dev)
local_r = torch.empty(B, dtype=torch.float64, device=dev)
for step in range(4):                             # four steps: both parities of the double buffer, twice
    tg, fl, stp = peer.begin()
    rc = L.icr_compute_rewards_gather(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, 224, 224,
                                      C.byref(prm), C.c_void_p(local_r.data_ptr()), C.c_void_p(tg.data_ptr()), C.c_void_p(fl.data_ptr()),
                                      C.c_ulonglong(stp), world, rank, ws.handle, st)
    assert rc == 0, L.icr_last_error()
    ema_p = reward.EMABaseline()
    reward.ema_update_gathered(ema_p, peer)       # waits on the device for every rank's flag of this step
    got = peer.gathered().clone()
    want = sharding.gather_rewards(local_r, B * world)
    assert torch.equal(got, want), (step, got, want)
    assert all(stp <= f <= stp + 1 for f in peer.flags.tolist())          # a rank may already be one step ahead
    ema_n = reward.EMABaseline()
    assert float(ema_p.value) == float(ema_n.update(want))
# the whole batch on one GPU gives the same vector, hence the same EMA value
full = reward.compute_rewards_batched(*(torch.from_numpy(a).to(dev) for a in (images, labels, sizes)))
assert torch.equal(full, got)
e1, e2 = reward.EMABaseline(), reward.EMABaseline()
assert float(e1.update(got)) == float(e2.update(full))
if rank == 0:
    print("PEER_GATHER_OK")
dist.destroy_process_group()
'''


@pytest.mark.gpu
def test_peer_store_gather_matches_nccl_all_gather(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29577", str(script)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "PEER_GATHER_OK" in out.stdout, out.stderr[-3000:]


NCCL_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch, torch.distributed as dist
from image_compression_cpp_b200 import reward, sharding, synth
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B = 16
images, labels, sizes = synth.make_batch(2, B=B)
lo, hi = sharding.shard_range(B, world, rank)
mine = reward.compute_rewards_batched(*(torch.from_numpy(np.ascontiguousarray(a[lo:hi])).to(dev) for a in (images, labels, sizes)))
gathered = sharding.gather_rewards(mine, B)
full = reward.compute_rewards_batched(*(torch.from_numpy(a).to(dev) for a in (images, labels, sizes)))
assert torch.equal(gathered, full), (gathered, full)          # bit for bit: same kernels, same per-image order
e1, e2 = reward.EMABaseline(), reward.EMABaseline()
assert float(e1.update(gathered)) == float(e2.update(full))
if rank == 0:
    print("NCCL_GATHER_OK")
dist.destroy_process_group()
'''


@pytest.mark.gpu
def test_nccl_gathered_rewards_equal_single_gpu_rewards(tmp_path):
    """Two ranks over NCCL: the all-gathered reward vector is bit-identical to the whole batch on one GPU."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(NCCL_WORKER % {"root": ROOT})
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29578", str(script)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "NCCL_GATHER_OK" in out.stdout, out.stderr[-3000:]


@pytest.mark.gpu
def test_two_devices_in_one_process():
    """The path on cuda:0 and then on cuda:1 from ONE process (the reference keeps 16 device slots per thread,
    src/png_size_estimator.cu:466-469): kernel attributes and grid sizes are per device."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import numpy as np
    from image_compression_cpp_b200 import reward, synth
    images, labels, sizes = synth.make_batch(2, B=4)
    got = []
    for d in (0, 1, 0):
        dev = torch.device("cuda", d)
        got.append(reward.compute_rewards_batched(*(torch.from_numpy(a).to(dev) for a in (images, labels, sizes))).cpu().numpy())
    assert np.array_equal(got[0], got[1]) and np.array_equal(got[0], got[2])
