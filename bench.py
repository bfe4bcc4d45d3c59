#!/usr/bin/env python
"""bench.py -- images/sec through the compression-reward path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step is one pass of the path over one batch of synthetic input.  Per GPU the batch is the
BASELINE config-5 shard: 512 images of 224x224 RGB (f32 planes, bytes 0..255) with BSP-rectangle
label maps (about 70 segments each, labels 7*id+3); N GPUs process N*512 images (weak scaling;
N=8 is config 5's 4096).  Inputs of one step (411 MB per GPU) exceed the 126 MB L2, so no flush is
needed between iterations.  `value` is timed with CUDA events on the launching stream with inputs
resident in HBM; `e2e` is the same metric through the C ABI's host-buffer entry
(icr_compute_rewards_host: pinned host buffers, H2D and D2H inside the timed region).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PER_GPU = int(os.environ.get("ICR_BENCH_PER_GPU", "512"))      # images per GPU and step (override only for experiments)
POOL = 64           # distinct synthetic images per rank; the shard tiles them (addresses differ, so no cache reuse)
H = W = 224
METRIC = "images/sec (slice+PNG-size reward)"


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.th.join(timeout=2)

    def summary(self):
        sm = [int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit()]
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def make_shard(rank):
    from image_compression_cpp_b200 import synth
    images, labels, sizes = synth.make_batch(5, B=POOL, seed0=5000 + rank * PER_GPU)
    reps = PER_GPU // POOL
    return np.tile(images, (reps, 1, 1, 1)), np.tile(labels, (reps, 1, 1)), np.tile(sizes, reps)


def bytes_per_image(K):
    # f32 planes 12 B/px + int32 labels 4 B/px; per segment count 4 + bbox 16 + size 8; size in 8 + reward out 8
    return H * W * 16 + K * 28 + 16


# --------------------------------------------------------------------------------------------
# CPU baseline: the reference's CPU slice path (src/image_slicer.cpp:48-74,98-108) + cv::imencode PNG level 4
# (src/image_writer.cpp:7, include/configuration.h:11), one task per image on all host cores.
def _cpu_one(args):
    import cv2
    cv2.setNumThreads(1)
    rgb_u8, lab = args
    bgra = np.concatenate([rgb_u8, np.full(rgb_u8.shape[:2] + (1,), 255, np.uint8)], axis=2)
    total = 0
    for label in np.unique(lab):
        mask = lab == label                                      # full-image pass per label (:51)
        ys, xs = np.nonzero(mask)                                # bbox scan (:15-41)
        y0, y1, x0, x1 = ys.min(), ys.max() + 1, xs.min(), xs.max() + 1
        out = np.zeros((y1 - y0, x1 - x0, 4), np.uint8)          # crop + masked copy onto zeros (:59-66)
        m = mask[y0:y1, x0:x1]
        out[m] = bgra[y0:y1, x0:x1][m]
        ok, buf = cv2.imencode(".png", out, [cv2.IMWRITE_PNG_COMPRESSION, 4])
        total += len(buf)
    return total


def cpu_baseline(images, labels, n_images, budget_s=20.0):
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    rgb = np.clip(np.rint(images[:n_images]), 0, 255).astype(np.uint8).transpose(0, 2, 3, 1)
    tasks = [(np.ascontiguousarray(rgb[i]), labels[i]) for i in range(n_images)]
    with mp.get_context("fork").Pool(cores) as pool:
        pool.map(_cpu_one, tasks[:cores])                        # warm the workers
        t0 = time.perf_counter()
        done = 0
        while True:
            pool.map(_cpu_one, tasks)
            done += len(tasks)
            if time.perf_counter() - t0 > budget_s / 2:
                break
        dt = time.perf_counter() - t0
    return {"value": done / dt, "unit": "images/sec", "cores": cores, "kind": "port",
            "sample": f"{done} images of the same workload ({n_images} distinct), numpy slice_image + cv2.imencode PNG level 4, "
                      f"one process per core, {dt:.1f} s"}


# --------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference arm: the UNMODIFIED reference sources (oracle/_ref/libref_reward.so, built by
    oracle/build_ref.sh from /root/reference/src/*.cu) through compute_rewards_batched, on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    images, labels, sizes = make_shard(0)
    n = PER_GPU                                                  # the same 512 images per step as the b200 arm (about 1.7 s per step: ~900 stream ops per image)
    path = os.path.join(ROOT, "oracle", "_ref", "libref_reward.so")
    line = {"impl": "reference", "metric": METRIC, "unit": "images/sec", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "cfg5 shard: 512 x 224x224 RGB f32 planes per GPU, BSP-rectangle int32 label maps (labels 7*id+3)", "per_gpu_batch": n, "global_batch": n}}
    if os.path.exists(path) and torch.cuda.is_available():
        L = C.CDLL(path)
        L.ref_last_error.restype = C.c_char_p
        dev = torch.device("cuda:0")
        ti = torch.from_numpy(images[:n]).to(dev)
        tl = torch.from_numpy(labels[:n]).to(dev)
        ts = torch.from_numpy(sizes[:n]).to(dev)
        out = torch.empty(n, dtype=torch.float64, device=dev)

        def step():
            rc = L.ref_compute_rewards(C.c_void_p(ti.data_ptr()), 1, C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()),
                                       n, H, W, 0, 1, 4, C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1),
                                       C.c_double(9.308622), 1, C.c_double(0.5), C.c_void_p(out.data_ptr()))
            if rc:
                raise RuntimeError(L.ref_last_error().decode())
        for _ in range(max(args.warmup, 1)):
            step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()                                               # the shim synchronises, as the reference's D2H does
        dt = time.perf_counter() - t0
        v = n * args.steps / dt
        line.update(value=v, ms_per_step=1e3 * dt / args.steps,
                    cpu_baseline={"value": v, "unit": "images/sec", "cores": 1, "kind": "reference",
                                  "sample": f"{n} images per step; the reference's own CUDA sources compiled unchanged for sm_100a, "
                                            "driven by one host thread on the same B200 (the reference has no CPU estimator)"})
    else:
        from oracle import oracle as orc
        t0 = time.perf_counter()
        for _ in range(args.steps):
            orc.compute_rewards(images[:n], labels[:n], sizes[:n])
        dt = time.perf_counter() - t0
        v = n * args.steps / dt
        line.update(value=v, ms_per_step=1e3 * dt / args.steps,
                    cpu_baseline={"value": v, "unit": "images/sec", "cores": 1, "kind": "port",
                                  "sample": f"{n} images per step, scalar C restatement (oracle/reward_oracle.c); oracle/_ref not available"})
    line["e2e"] = {"value": line["value"], "unit": "images/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    print(json.dumps(line), file=args.out, flush=True)


def bind_to_gpu_numa_node(local):
    """One process per GPU: run on the cores of the NUMA node the GPU hangs off, so that the pinned host buffers of the
    host-buffer entry are first touched in that node's memory and the H2D copies do not cross the socket link.
    Best effort (sysfs + sched_setaffinity); returns the node or None."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


# --------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    # stdout carries exactly one JSON line: whatever libraries write to file descriptor 1 on the way (NCCL prints its
    # version banner there when NCCL_DEBUG is set) goes to stderr instead, and the line is written to the real stdout.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args.out = real_stdout
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from image_compression_cpp_b200 import reward, sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    numa = bind_to_gpu_numa_node(local)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = reward.lib()

    images, labels, sizes = make_shard(rank)
    B = PER_GPU
    Kmean = float(np.mean([len(np.unique(l)) for l in labels[:POOL]]))
    ti, tl, ts = torch.from_numpy(images).to(dev), torch.from_numpy(labels).to(dev), torch.from_numpy(sizes).to(dev)
    ws = reward.Workspace(dev, B, H, W, Kmax=1024)
    prm = reward.make_params()
    rewards = torch.empty(B, dtype=torch.float64, device=dev)
    gathered = torch.empty(B * world, dtype=torch.float64, device=dev)
    sum_count = torch.empty(2, dtype=torch.float64, device=dev)
    ema = reward.EMABaseline()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    # N > 1: the path's only exchange is that every rank receives all per-image rewards: one NCCL all-gather per step.
    # ICR_BENCH_EXCHANGE=peer selects the fused variant instead (the reward kernel stores every reward straight into all
    # ranks' gathered vectors through peer memory and a step needs one device-side barrier); measured equal at N=2
    # (0.2945 ms per step both ways: the all-gather already runs under the next step's first kernel), so the plain
    # collective stays the default.
    peer = None
    if world > 1 and os.environ.get("ICR_BENCH_EXCHANGE", "nccl") == "peer":
        ok = torch.ones(1, device=dev)
        try:
            peer = sharding.PeerGather(B, dev)
        except Exception as e:                      # noqa: BLE001 -- any failure means "not available here"
            print(f"[bench] rank {rank}: peer gather unavailable ({type(e).__name__}: {e}); using the NCCL all-gather", file=sys.stderr)
            ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() == 0:
            peer = None
    exchange = "none" if world == 1 else ("peer stores over NVLink + flags (no collective)" if peer else
                                             "NCCL all-gather on a side stream under the next step's kernels")
    if world > 1 and os.environ.get("ICR_BENCH_EXCHANGE") == "none":
        exchange = "NOTHING (diagnostic run, not a valid bench line)"

    # The exchange of step s and its EMA run on a side stream while the main stream already computes step s+1
    # (sharding.PipelinedGather): ranks do not wait for each other inside every step, only at the end of the timed region.
    pipe = sharding.PipelinedGather(B, dev) if world > 1 and peer is None and os.environ.get("ICR_BENCH_EXCHANGE") != "none" else None

    def drain():
        if pipe is not None:
            pipe.drain()

    def step_pipelined():
        buf = pipe.next_buffer()
        rc = L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, H, W,
                                   C.byref(prm), C.c_void_p(buf.data_ptr()), ws.handle, stream)
        if rc:
            raise RuntimeError(L.icr_last_error().decode())
        pipe.submit(ema.update)

    def step():
        if pipe is not None:
            return step_pipelined()
        if peer is not None:
            tg, fl, stp = peer.begin()
            rc = L.icr_compute_rewards_gather(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, H, W,
                                              C.byref(prm), C.c_void_p(rewards.data_ptr()), C.c_void_p(tg.data_ptr()),
                                              C.c_void_p(fl.data_ptr()), C.c_ulonglong(stp), world, rank, ws.handle, stream)
        else:
            rc = L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, H, W,
                                       C.byref(prm), C.c_void_p(rewards.data_ptr()), ws.handle, stream)
        if rc:
            raise RuntimeError(L.icr_last_error().decode())
        if peer is not None:
            reward.ema_update_gathered(ema, peer)
        elif world > 1 and os.environ.get("ICR_BENCH_EXCHANGE") != "none":      # "none": diagnosis only, prints an invalid line
            dist.all_gather_into_tensor(gathered, rewards)
            ema.update(gathered)
        else:
            ema.update(rewards)

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        step()
    ws.check()
    sync()
    L.icr_launch_count(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        sync()
        e0.record()
        for _ in range(args.steps):
            step()
        drain()
        e1.record()
        sync()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = int(L.icr_launch_count(1))
    value = B * world * args.steps / (ms * 1e-3)

    # ---- e2e: host buffers through the C ABI, H2D + D2H inside the timed region
    pi, pl, ps = torch.from_numpy(images).pin_memory(), torch.from_numpy(labels).pin_memory(), torch.from_numpy(sizes).pin_memory()
    pr = torch.empty(B, dtype=torch.float64).pin_memory()

    def step_host():
        rc = L.icr_compute_rewards_host(C.c_void_p(pi.data_ptr()), C.c_void_p(pl.data_ptr()), 0, C.c_void_p(ps.data_ptr()), B, 1, H, W,
                                        C.byref(prm), C.c_void_p(pr.data_ptr()), ws.handle, stream)
        if rc:
            raise RuntimeError(L.icr_last_error().decode())
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(3):
        step_host()
    sync()
    e0.record()
    for _ in range(e2e_steps):
        step_host()
    e1.record()
    sync()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))
    e2e_value = B * world * e2e_steps / (e2e_ms * 1e-3)
    L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, H, W,
                          C.byref(prm), C.c_void_p(rewards.data_ptr()), ws.handle, stream)        # untimed: the device entry's answer
    assert np.array_equal(pr.numpy(), rewards.cpu().numpy()), "host-buffer entry and device entry disagree"
    # what the link alone takes for the same pinned buffers: e2e cannot go below this
    scratch_i, scratch_l = torch.empty_like(ti), torch.empty_like(tl)
    for _ in range(2):
        scratch_i.copy_(pi, non_blocking=True); scratch_l.copy_(pl, non_blocking=True)
    sync()
    e0.record()
    for _ in range(5):
        scratch_i.copy_(pi, non_blocking=True); scratch_l.copy_(pl, non_blocking=True)
    e1.record()
    sync()
    copy_ms = max_over_ranks(e0.elapsed_time(e1)) / 5
    del scratch_i, scratch_l

    # ---- roofline of the dominant kernel: per-kernel device times from events between the launches
    L.icr_workspace_set_profile(ws.handle, 1)
    kms = np.zeros((args.steps, 4), np.float32)
    for i in range(args.steps):
        L.icr_compute_rewards(C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()), B, 1, H, W,
                              C.byref(prm), C.c_void_p(rewards.data_ptr()), ws.handle, stream)
        L.icr_profile_read(ws.handle, kms[i].ctypes.data_as(C.c_void_p))
    L.icr_workspace_set_profile(ws.handle, 0)
    kavg = kms.mean(0).astype(np.float64)
    names = ["prep_kernel", "narrow_kernel", "big_kernel", "reserved"]
    top = int(np.argmax(kavg))
    peak, peak_src = peaks()
    alg_bytes = bytes_per_image(Kmean) * B
    achieved = float(alg_bytes / (kavg[top] * 1e-3) / 1e9)
    path_gbs = float(alg_bytes / (ms / args.steps * 1e-3) / 1e9)

    traffic = None
    try:                          # dram bytes per launch of the same kernel from the committed ncu --set full capture
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            traffic = json.load(f).get(names[top])
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": "images/sec", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": "cfg5 shard: 512 x 224x224 RGB f32 planes per GPU, BSP-rectangle int32 label maps (labels 7*id+3)",
                   "per_gpu_batch": B, "global_batch": B * world, "mean_segments_per_image": Kmean, "l2": "inputs (411 MB/step/GPU) exceed L2; no flush",
                   "parallelism": f"dp{world}: images sharded, rewards exchanged by {exchange}, EMA on the gathered vector" if world > 1 else "single GPU"},
        "e2e": {"value": e2e_value, "unit": "images/sec", "h2d_bytes_per_step": int(images.nbytes + labels.nbytes + sizes.nbytes),
                "d2h_bytes_per_step": int(pr.numel() * 8), "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                "h2d_copy_alone_ms": copy_ms, "numa_node": numa,
                "api": "icr_compute_rewards_host (C ABI, pinned host buffers; the batch is copied in four pieces on a second stream "
                       "under the kernels of the piece before)"},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "kernel": names[top], "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": "profiles/r01_traffic.json (ncu --set full, per launch)" if traffic else None,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": float(alg_bytes),
                     "kernel_ms": {n: float(k) for n, k in zip(names, kavg)}, "whole_path_gbs": path_gbs, "whole_path_frac": path_gbs / peak},
        "clocks": clk.summary(),
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(images, labels, n_images=POOL)
    if rank == 0:
        print(json.dumps(line), file=args.out, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
