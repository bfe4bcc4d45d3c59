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


@pytest.mark.gpu
def test_cabi_exchange_on_a_one_rank_communicator():
    """icr_exchange_rewards on one GPU (a communicator of one rank, made through the C ABI): the all-gather is a copy, the
    all-reduce leaves this rank's {sum, count}; the C++ EMABaseline::update_global(rewards, comm) on top gives update()'s
    value.  This is the leg of the exchange the single-GPU driver box can see; the two-rank test below skips there."""
    import ctypes as C
    import numpy as np
    from image_compression_cpp_b200 import reward, sharding
    dev = torch.device("cuda:0")
    ex = sharding.NcclExchange(dev)
    assert ex.world == 1
    ranks = C.c_int(0)
    assert reward.lib().icr_nccl_comm_ranks(ex.comm, C.byref(ranks)) == 0 and ranks.value == 1
    rng = np.random.default_rng(3)
    for n in (1, 37, 4096):
        r = torch.from_numpy(rng.normal(0.3, 0.2, n)).to(dev)
        g = torch.zeros(n, dtype=torch.float64, device=dev)
        ex(g, r)
        torch.cuda.synchronize()
        assert torch.equal(g, r)
        sc = ex.sum_count.cpu().numpy()
        assert sc[1] == n and abs(sc[0] - r.sum().item()) <= 1e-12 * max(1.0, abs(r.sum().item()))
    # bad arguments
    L = reward.lib()
    assert L.icr_exchange_rewards(None, C.c_void_p(r.data_ptr()), n, C.c_void_p(g.data_ptr()), None, None) != 0
    assert L.icr_exchange_rewards(ex.comm, C.c_void_p(r.data_ptr()), n, None, None, None) != 0
    assert "nothing to exchange" in L.icr_last_error().decode()
    # the C++ layer on top
    host_lib = os.path.join(ROOT, "image-compression-cpp_b200", "libicr_host.so")
    H = C.CDLL(host_lib)
    H.icrh_last_error.restype = C.c_char_p
    value = C.c_double(0)
    rc = H.icrh_ema_update_global_nccl(C.c_void_p(r.data_ptr()), n, 0, ex.comm, 1, C.byref(value), C.c_void_p(g.data_ptr()))
    assert rc == 0, H.icrh_last_error()
    assert value.value == float(reward.EMABaseline().update(r)) and torch.equal(g, r)
    ex.close()


CABI_WORKER = r'''
import ctypes as C, os, sys
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
ex = sharding.NcclExchange(dev)
gathered = torch.zeros(B, dtype=torch.float64, device=dev)
for _ in range(3):
    ex(gathered, mine)
torch.cuda.synchronize()
full = reward.compute_rewards_batched(*(torch.from_numpy(a).to(dev) for a in (images, labels, sizes)))
assert torch.equal(gathered, full), (gathered, full)
sc = ex.sum_count.cpu().numpy()
assert sc[1] == B and abs(sc[0] - full.sum().item()) <= 1e-12 * abs(full.sum().item())
# pipelined, as bench.py --exchange cabi runs it, and the C++ EMABaseline::update_global(rewards, comm)
pipe = sharding.PipelinedGather(hi - lo, dev, gather=ex)
emas = []
for step in range(4):
    buf = pipe.next_buffer(); buf.copy_(mine)
    e = reward.EMABaseline(); emas.append(e); pipe.submit(e.update)
pipe.drain(); torch.cuda.synchronize()
want = float(reward.EMABaseline().update(full))
assert all(float(e.value) == want for e in emas)
H = C.CDLL(os.path.join(%(root)r, "image-compression-cpp_b200", "libicr_host.so")); H.icrh_last_error.restype = C.c_char_p
value = C.c_double(0); g2 = torch.zeros(B, dtype=torch.float64, device=dev)
rc = H.icrh_ema_update_global_nccl(C.c_void_p(mine.data_ptr()), hi - lo, local, ex.comm, world, C.byref(value), C.c_void_p(g2.data_ptr()))
assert rc == 0, H.icrh_last_error()
assert value.value == want and torch.equal(g2, full)
ex.close()
if rank == 0:
    print("CABI_EXCHANGE_OK")
dist.destroy_process_group()
'''


@pytest.mark.gpu
def test_cabi_exchange_two_ranks(tmp_path):
    """Two ranks: icr_exchange_rewards (ncclAllGather + ncclAllReduce in one group, bound from the process's libnccl at run time)
    delivers the whole batch's rewards bit for bit, in line and through the pipelined side stream, and the C++
    EMABaseline::update_global(rewards, comm) built on it returns the single-GPU baseline."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(CABI_WORKER % {"root": ROOT})
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29579", str(script)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "CABI_EXCHANGE_OK" in out.stdout, out.stderr[-3000:]
