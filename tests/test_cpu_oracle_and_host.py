"""CPU suite (-m "not gpu"): the oracle against the reference-generated golden vectors and its own invariants,
the host logic, and that the C-ABI library loads and exports every symbol include/icr_b200.h declares."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from image_compression_cpp_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_cases.npz")


def _cases():
    if not os.path.exists(GOLDEN):
        return {}
    z = np.load(GOLDEN)
    names = sorted({k.split("/")[0] for k in z.files})
    return {n: {k.split("/")[1]: z[k] for k in z.files if k.startswith(n + "/")} for n in names}


CASES = _cases()


def _prm(orc, p):
    return orc.params(int(p[0]), int(p[1]), float(p[2]), float(p[3]), float(p[4]), float(p[5]), bool(p[6]), float(p[7]))


def test_golden_fixture_present():
    assert CASES, "tests/golden/reference_cases.npz is missing: run tests/golden/make_golden.py on the GPU box"


@pytest.mark.parametrize("name", sorted(CASES) or ["none"])
def test_oracle_matches_reference_outputs(orc, name):
    """The pin: oracle/reward_oracle.c against outputs of the reference's own CUDA sources run on a B200."""
    if not CASES:
        pytest.skip("no golden fixture")
    c = CASES[name]
    prm = _prm(orc, c["params"])
    r, det = orc.compute_rewards(c["images"], c["labels"], c["sizes"], prm, details=True)
    assert np.allclose(r, c["rewards"], rtol=1e-11, atol=0), (name, r, c["rewards"])
    K = len(c["counts0"])
    assert det["K"][0] == K
    assert np.array_equal(det["counts"][0, :K], c["counts0"])
    assert np.array_equal(det["bboxes"][0, :K], c["bboxes0"])
    assert np.allclose(det["seg_sizes"][0, :K], c["seg_sizes0"], rtol=1e-12, atol=0)


def test_oracle_slice_equals_filter_on_cropped_slice(orc):
    """Rule P: masking neighbours inside the bbox == PNG-filtering the materialised slice (what slice_image produces)."""
    img = synth.image_u8(5, 40, 48, 1)
    rgba = np.concatenate([img, np.full((40, 48, 1), 255, np.uint8)], 2)
    lab = synth.bfs_partition(40, 48, 5, seed=3).astype(np.int64)
    comp, uniq = orc.compact_labels(lab)
    counts, bboxes = orc.segment_stats(comp, len(uniq))
    assert counts.sum() == 40 * 48
    for k in range(len(uniq)):
        x0, y0, x1, y1 = bboxes[k]
        w, h = x1 - x0 + 1, y1 - y0 + 1
        sl = orc.slice_segment(rgba, comp, k, x0, y0, w, h)
        m = comp[y0:y1 + 1, x0:x1 + 1] == k
        assert np.array_equal(sl[m], rgba[y0:y1 + 1, x0:x1 + 1][m]) and not sl[~m].any()
        s, tab = orc.estimate_segment(rgba, comp, k, x0, y0, w, h, debug=True)
        # histogram rows each sum to w*h; the None-filter cost of a row is sum |int8(byte)|
        assert np.array_equal(tab["hist"].sum(1), np.full(4, w * h))
        none_cost = np.abs(sl.astype(np.int8).astype(np.int32)).reshape(h, -1).sum(1)
        assert np.array_equal(tab["row_costs"][:, 0], none_cost)
        # ties go to the lowest filter id
        assert np.array_equal(tab["filter_id"], np.argmin(tab["row_costs"], 1))
        assert tab["match"].tolist() == [0, 0, 0]         # N <= 65536*3: chunks shorter than L_min


def test_oracle_rule_m_counts_runs_inside_chunks(orc):
    H = W = 224
    rgba = np.zeros((H, W, 4), np.uint8)
    rgba[..., 3] = 255
    rgba[..., :3] = 77                                     # monochrome: Sub residuals are all zero after column 0
    comp = np.zeros((H, W), np.int64)
    s, tab = orc.estimate_segment(rgba, comp, 0, 0, 0, W, H, debug=True)
    N = H * W * 4
    chunk = (N + 65535) // 65536
    assert chunk == 4
    res = tab["residuals"].reshape(-1, 4)
    runs = (res == res[:, :1]).all(1).sum()                # a 4-byte chunk matches iff its 4 bytes are equal
    assert tab["match"].tolist() == [4 * runs, runs, 4 * runs]


def test_oracle_ema_rule(orc):
    v, init = orc.ema_update(0.0, False, 0.99, [1.0, 2.0, 3.0])
    assert init and v == 2.0
    v2, _ = orc.ema_update(v, True, 0.99, [4.0])
    assert abs(v2 - (2.0 * 0.99 + 4.0 * 0.01)) < 1e-15


def test_c_abi_exports_every_declared_symbol(icr):
    hdr = open(os.path.join(ROOT, "include", "icr_b200.h")).read()
    declared = set(re.findall(r"\b(icr_[a-z0-9_]+)\s*\(", hdr)) - {"icr_workspace", "icr_params", "icr_stream"}
    L = ctypes.CDLL(icr.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert set(icr.SYMBOLS) <= declared
    out = subprocess.check_output(["nm", "-D", "--defined-only", icr.LIB_PATH], text=True)
    assert "torch" not in out and "at::" not in out        # the boundary is torch-free


def test_c_abi_argument_errors_without_a_gpu(icr):
    L = icr.lib()
    p = icr.make_params()
    assert L.icr_compute_rewards(None, None, 0, None, 1, 1, 8, 8, ctypes.byref(p), None, None, None) == 1
    assert b"null ptr" in L.icr_last_error()
    assert L.icr_estimate_png_size_segment(None, 8, 8, 4, None, None, 0, 0, 0, 8, 8, 1, 4, ctypes.c_float(0), ctypes.c_float(0),
                                           ctypes.c_float(0), ctypes.c_double(0), 1, None, None, None, None, None, None) == 1
    assert 4 * 64 * 36 < L.icr_workspace_bytes(4, 224, 224, 64) < 1 << 20        # no per-pixel plane in the workspace
    assert L.icr_label_capacity(224, 224) == 65536 and L.icr_label_capacity(512, 512) == 512 * 512
    # the entry points added in round 2 refuse bad arguments before they touch a device
    assert L.icr_estimate_png_size_segment_noalloc(None, 8, 8, 4, None, None, 0, 0, 0, 8, 8, 1, 4, ctypes.c_float(0), ctypes.c_float(0),
                                                   ctypes.c_float(0), ctypes.c_double(0), 1, None, None, None, None, None, None) == 1
    assert L.icr_compute_rewards_host_u8(None, 7, None, 0, None, 1, 1, 8, 8, ctypes.byref(p), None, None, None) == 1
    assert b"image_layout" in L.icr_last_error()
    assert L.icr_compute_rewards_host_u8(None, 0, None, 0, None, 1, 1, 8, 8, ctypes.byref(p), None, None, None) == 1
    assert L.icr_pack_node_labels(None, None, 2, ctypes.c_int64(64), None, None) == 1
    assert L.icr_exchange_rewards(None, None, 4, None, None, None) == 1
    assert L.icr_nccl_comm_create(None, None, 2, 0) == 1 and L.icr_nccl_comm_ranks(None, None) == 1
    assert L.icr_nccl_comm_destroy(None) == 0
    q = icr.IcrParams()
    L.icr_default_params(ctypes.byref(q))
    assert (q.min_pixels, q.L_min, q.adaptive_filter, q.lambda_) == (1, 4, 1, 0.5) and abs(q.overhead_base - 9.308622) < 1e-15


def test_host_mirror_rejects_what_the_reference_rejects(icr):
    im = torch.zeros(2, 3, 8, 8)
    lab = torch.zeros(2, 8, 8, dtype=torch.int32)
    sz = torch.ones(2, dtype=torch.float64)
    with pytest.raises(RuntimeError, match="labels must be CUDA"):       # src/compute_rewards.cu:72 -- and no CPU fallback
        icr.compute_rewards_batched(im, lab, sz)
    with pytest.raises(RuntimeError, match=r"\[B,3,H,W\]"):
        icr.compute_rewards_batched(im[:, :1], lab, sz)
    with pytest.raises(RuntimeError, match="float32"):
        icr.compute_rewards_batched(im.double(), lab, sz)
    with pytest.raises(RuntimeError, match="labels must be CUDA"):
        icr.compute_counts_bboxes_from_compact_labels_cuda(lab[0].long(), torch.zeros(1, dtype=torch.int32), torch.zeros(1, 4, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="CUDA"):
        icr.labels_from_rama(torch.zeros(2, 64, dtype=torch.int32), 8, 8)


def test_synth_is_seeded_and_shaped():
    a = synth.make_batch(2, B=3)
    b = synth.make_batch(2, B=3)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert a[0].shape == (3, 3, 224, 224) and a[0].dtype == np.float32 and a[1].dtype == np.int32
    assert a[0].min() >= 0 and a[0].max() <= 255 and np.array_equal(a[0], np.rint(a[0]))
    K = [len(np.unique(l)) for l in a[1]]
    assert all(60 <= k <= 90 for k in K)
    assert len(np.unique(synth.voronoi_partition(512, 512, seed=3))) == 1024
    im, lab, _ = synth.make_multi_batch(B=2, S=3, H=32, W=40)
    assert lab.shape == (2, 3, 32, 40)


def test_two_rank_gloo_exchange_matches_single_process(orc, tmp_path):
    """world_size 2 on CPU (gloo): sharding + all-gather of rewards + all-reduce of the EMA statistics."""
    script = tmp_path / "w.py"
    script.write_text(f"""
import os, sys
sys.path.insert(0, {ROOT!r})
import numpy as np, torch, torch.distributed as dist
from image_compression_cpp_b200 import sharding
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
B = 7
allr = torch.arange(B, dtype=torch.float64) * 0.25 - 0.5
lo, hi = sharding.shard_range(B, world, rank)
local = allr[lo:hi].clone()
sc = torch.tensor([float(local.sum()), float(local.numel())], dtype=torch.float64)
g, tot = sharding.exchange(local, sc, B)
assert torch.equal(g, allr), (g, allr)
assert tot.tolist() == [float(allr.sum()), float(B)]
assert torch.equal(sharding.gather_rewards(local, B), allr)
if rank == 0:
    print("OK", float(tot[0] / tot[1]))
dist.destroy_process_group()
""")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29571")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29571", str(script)], capture_output=True, text=True, env=env, timeout=240)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "OK" in out.stdout
    mean = float(out.stdout.split("OK")[1].split()[0])
    v, _ = orc.ema_update(0.0, False, 0.99, np.arange(7) * 0.25 - 0.5)
    assert abs(mean - v) < 1e-15


def test_exchange_classes_are_importable_without_a_gpu():
    """PipelinedGather / PeerGather need CUDA to construct; their module must still import on a CPU box."""
    from image_compression_cpp_b200 import sharding
    assert callable(sharding.PipelinedGather) and callable(sharding.PeerGather) and sharding.shard_range(10, 4, 3) == (8, 10)


def test_bench_workload_sharding():
    """bench.py's workload builder on CPU: strong scaling cuts the config's global batch into equal contiguous shards with
    their own synthetic images, weak scaling keeps the per-GPU batch, and the algorithmic bytes follow BASELINE.md section 3."""
    import bench
    a = bench.Workload(2, 0, 2, "strong")
    b = bench.Workload(2, 1, 2, "strong")
    assert a.B == b.B == 32 and a.global_batch == 64
    assert a.images.shape == (32, 3, 224, 224) and a.labels.shape == (32, 224, 224) and a.labels.dtype == np.int32
    assert not np.array_equal(a.images[0], b.images[0])                       # rank 1 holds other images than rank 0
    whole = bench.Workload(2, 0, 1, "strong")
    assert whole.B == 64 and np.array_equal(whole.images[:32], a.images) and np.array_equal(whole.labels[32:], b.labels)
    assert abs(whole.bytes_per_image() - (224 * 224 * 16 + whole.Kmean * 28 + 16)) < 1e-6
    with pytest.raises(SystemExit):
        bench.Workload(2, 0, 3, "strong")                                      # 64 images do not split over 3 GPUs
    assert bench.CONFIGS[5]["G"] == 4096 and bench.CONFIGS[3]["H"] == 512 and bench.CONFIGS[4]["S"] == 8
