#!/usr/bin/env python
"""bench.py -- images/sec through the compression-reward path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config 5|2|3|4] [--scaling strong|weak]

A step is one pass of the path over one batch of synthetic input.  The default workload is BASELINE config 5 as written:
4096 images of 224x224 RGB (f32 planes, bytes 0..255) with BSP-rectangle label maps (about 70 segments each, labels
7*id+3), sharded over the N GPUs (strong scaling: 4096/N images per GPU; N=1 holds all 4096, 3.3 GB).  --scaling weak
keeps 512 images per GPU instead (last round's line).  --config 3 (256 x 512x512, Voronoi maps of 1024 segments) and
--config 4 (256 x 224x224 with 8 label maps per image) and --config 2 (64 x 224x224, the batch of one training step; it
fits the L2, so every step is timed on its own with the L2 flushed before it) print the same line for the other configs.
Inputs of one step exceed the 126 MB L2 for configs 5, 3 and 4, so no flush is needed between their iterations.
`value` is timed with CUDA events on the launching stream with inputs resident in HBM; `e2e` is the same metric through
the C ABI's host-buffer entry (icr_compute_rewards_host: pinned host buffers, H2D and D2H inside the timed region).
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

WEAK_PER_GPU = int(os.environ.get("ICR_BENCH_PER_GPU", "512"))   # images per GPU and step under --scaling weak
METRIC = "images/sec (slice+PNG-size reward)"
# BASELINE.json configs: global batch, label maps per image, shape, table rows, distinct synthetic images per rank (the
# batch tiles them: addresses differ, so nothing is reused from cache)
CONFIGS = {
    2: dict(G=64, S=1, H=224, W=224, Kmax=1024, pool=64,
            name="cfg2: 64 x 224x224 RGB f32 planes, BSP-rectangle int32 label maps (~72 segments)"),
    5: dict(G=4096, S=1, H=224, W=224, Kmax=1024, pool=64,
            name="cfg5: 4096 x 224x224 RGB f32 planes, BSP-rectangle int32 label maps (labels 7*id+3)"),
    3: dict(G=256, S=1, H=512, W=512, Kmax=1024, pool=32,
            name="cfg3: 256 x 512x512 RGB f32 planes, Voronoi (SLIC-style) int32 label maps of 1024 segments"),
    4: dict(G=256, S=8, H=224, W=224, Kmax=512, pool=16,
            name="cfg4: 256 x 224x224 RGB f32 planes, 8 int32 label maps per image (BSP rectangles / region growing)"),
}


class Workload:
    """One rank's share of a config: numpy inputs + shape."""

    def __init__(self, cfg, rank, world, scaling):
        from image_compression_cpp_b200 import synth
        c = CONFIGS[cfg]
        self.cfg, self.S, self.H, self.W, self.Kmax, self.name = cfg, c["S"], c["H"], c["W"], c["Kmax"], c["name"]
        if scaling == "weak":
            self.B = WEAK_PER_GPU if cfg == 5 else c["G"]
        else:
            if c["G"] % world:
                raise SystemExit(f"config {cfg}: global batch {c['G']} does not split over {world} GPUs")
            self.B = c["G"] // world
        self.global_batch = self.B * world
        pool = min(c["pool"], self.B)
        first = rank * self.B                                       # index of this rank's first image in the global batch
        if cfg == 2:
            images, labels, sizes = synth.make_batch(2, B=pool, seed0=1000 + first)
        elif cfg == 5:
            images, labels, sizes = synth.make_batch(5, B=pool, seed0=5000 + first)
        elif cfg == 3:
            images, labels, sizes = synth.make_batch(3, B=pool, seed0=3000 + first)
        else:
            images, labels, sizes = synth.make_multi_batch(B=pool, S=self.S, seed0=4000 + first * self.S)
        self.pool = pool
        reps = -(-self.B // pool)
        self.images = np.ascontiguousarray(np.tile(images, (reps, 1, 1, 1))[:self.B])
        self.labels = np.ascontiguousarray(np.tile(labels, (reps,) + (1,) * (labels.ndim - 1))[:self.B])
        self.sizes = np.ascontiguousarray(np.tile(sizes, reps)[:self.B])
        maps = labels.reshape(-1, self.H, self.W)
        self.Kmean = float(np.mean([len(np.unique(m)) for m in maps[:64]]))

    def bytes_per_image(self):
        # f32 planes 12 B/px + int32 labels 4 B/px per label map; per segment count 4 + bbox 16 + size 8;
        # per map image_size in 8 + reward out 8
        return self.H * self.W * (12 + 4 * self.S) + self.S * (self.Kmean * 28 + 16)

    def step_bytes(self):
        return int(self.images.nbytes + self.labels.nbytes + self.sizes.nbytes)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock and throttle reasons of this rank's GPU, polled through NVML every few milliseconds by a thread while the
    timed region runs (the region lasts tens of milliseconds: nvidia-smi's own loop is too coarse for it; it stays as the
    fallback when the NVML binding is missing)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self.stop, self.th, self.proc, self.rows, self.source = threading.Event(), None, None, [], None

    def _nvml_loop(self, nv, h):
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        while not self.stop.is_set():
            try:
                self.sm.append(int(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                self.reasons.update(n for n, b in bits.items() if r & b)
            except Exception:
                pass
            self.stop.wait(0.004)

    def __enter__(self):
        try:
            import pynvml as nv
            import torch
            nv.nvmlInit()
            p = torch.cuda.get_device_properties(self.index)
            h = nv.nvmlDeviceGetHandleByPciBusId(("%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)).encode())
            self.max_mhz = int(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            nv.nvmlDeviceGetCurrentClocksEventReasons(h)
            self.th = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self.th.start()
            self.source = "nvml, 4 ms period"
            return self
        except Exception:
            self.th = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
            self.source = "nvidia-smi -lms 100"
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
        elif self.th:
            self.stop.set()
            self.th.join(timeout=2)

    def summary(self):
        if self.proc:
            self.sm = [int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit()]
            mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
            self.max_mhz = max(mx) if mx else None
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            self.reasons = {n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")}
        return {"sm_mhz": int(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.sm), "source": self.source}


# --------------------------------------------------------------------------------------------
# CPU baseline: the reference's CPU slice path (src/image_slicer.cpp:48-74,98-108) + cv::imencode PNG level 4
# (src/image_writer.cpp:7, include/configuration.h:11), one task per image on all host cores.
def _cpu_one(args):
    import cv2
    cv2.setNumThreads(1)
    rgb_u8, labs = args                                              # labs [S,H,W]: every label map of this image
    bgra = np.concatenate([rgb_u8, np.full(rgb_u8.shape[:2] + (1,), 255, np.uint8)], axis=2)
    totals = []
    for lab in labs:
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
        totals.append(total)
    return totals


def cpu_baseline(wl, budget_s=20.0):
    """Returns (cpu_baseline dict, encoded bytes per label map of the pool [pool*S])."""
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    n = wl.pool
    rgb = np.clip(np.rint(wl.images[:n]), 0, 255).astype(np.uint8).transpose(0, 2, 3, 1)
    labs = wl.labels[:n].reshape(n, wl.S, wl.H, wl.W)
    tasks = [(np.ascontiguousarray(rgb[i]), labs[i]) for i in range(n)]
    with mp.get_context("fork").Pool(cores) as pool:
        pool.map(_cpu_one, tasks[:cores])                            # warm the workers
        t0 = time.perf_counter()
        done = 0
        while True:
            totals = pool.map(_cpu_one, tasks)
            done += len(tasks)
            if time.perf_counter() - t0 > budget_s / 2:
                break
        dt = time.perf_counter() - t0
    what = "image" if wl.S == 1 else f"image ({wl.S} label maps each)"
    return ({"value": done / dt, "unit": "images/sec", "cores": cores, "kind": "port",
             "sample": f"{done} images of the same workload ({n} distinct), numpy slice_image + cv2.imencode PNG level 4 (a real "
                       f"deflate encoder: more work per byte than the size estimator, so this is the CPU path the reference "
                       f"replaces, not the same arithmetic), one task per {what}, one process per core, {dt:.1f} s"},
            np.asarray(totals, np.float64).reshape(-1))


# --------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference arm: the UNMODIFIED reference sources (oracle/_ref/libref_reward.so, built by
    oracle/build_ref.sh from /root/reference/src/*.cu) through compute_rewards_batched, on the b200 arm's config; each
    step is a bounded sample of that batch so that the run ends within minutes (about 300 stream operations per image
    and two host syncs: 3.4 ms per config-5 image)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    wl = Workload(args.config, 0, 1, "strong")
    n = min(wl.B, {5: 512, 3: 16, 4: 32, 2: 64}[args.config])    # images per step
    H, W, S = wl.H, wl.W, wl.S
    images = np.repeat(wl.images[:n], S, axis=0) if S > 1 else wl.images[:n]      # the reference's own formulation of S maps per image
    labels = wl.labels[:n].reshape(n * S, H, W)
    sizes = np.repeat(wl.sizes[:n], S)
    path = os.path.join(ROOT, "oracle", "_ref", "libref_reward.so")
    line = {"impl": "reference", "metric": METRIC, "unit": "images/sec", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl.name, "global_batch": CONFIGS[args.config]["G"], "images_per_reference_step": n,
                       "mean_segments_per_label_map": wl.Kmean}}
    if os.path.exists(path) and torch.cuda.is_available():
        L = C.CDLL(path)
        L.ref_last_error.restype = C.c_char_p
        dev = torch.device("cuda:0")
        ti, tl, ts = torch.from_numpy(images).to(dev), torch.from_numpy(labels).to(dev), torch.from_numpy(sizes).to(dev)
        out = torch.empty(n * S, dtype=torch.float64, device=dev)

        def step():
            rc = L.ref_compute_rewards(C.c_void_p(ti.data_ptr()), 1, C.c_void_p(tl.data_ptr()), 0, C.c_void_p(ts.data_ptr()),
                                       n * S, H, W, 0, 1, 4, C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1),
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
                                  "sample": f"{n} images of the config per step; the reference's own CUDA sources compiled unchanged for sm_100a, "
                                            "driven by one host thread on the same B200 (the reference has no CPU estimator)"})
    else:
        from oracle import oracle as orc
        t0 = time.perf_counter()
        for _ in range(args.steps):
            orc.compute_rewards(images, labels, sizes)
        dt = time.perf_counter() - t0
        v = n * args.steps / dt
        line.update(value=v, ms_per_step=1e3 * dt / args.steps,
                    cpu_baseline={"value": v, "unit": "images/sec", "cores": 1, "kind": "port",
                                  "sample": f"{n} images of the config per step, scalar C restatement (oracle/reward_oracle.c); oracle/_ref not available"})
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
    ap.add_argument("--config", type=int, default=5, choices=sorted(CONFIGS))
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
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

    wl = Workload(args.config, rank, world, args.scaling)
    B, S, H, W = wl.B, wl.S, wl.H, wl.W
    V = B * S                                                        # label maps = rewards per rank and step
    ti, tl, ts = torch.from_numpy(wl.images).to(dev), torch.from_numpy(wl.labels).to(dev), torch.from_numpy(wl.sizes).to(dev)
    ws = reward.Workspace(dev, V, H, W, Kmax=wl.Kmax)
    prm = reward.make_params()
    rewards = torch.empty(V, dtype=torch.float64, device=dev)
    gathered = torch.empty(V * world, dtype=torch.float64, device=dev)
    ema = reward.EMABaseline()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p_i, p_l, p_s = C.c_void_p(ti.data_ptr()), C.c_void_p(tl.data_ptr()), C.c_void_p(ts.data_ptr())

    def device_call(out):
        rc = L.icr_compute_rewards(p_i, p_l, 0, p_s, B, S, H, W, C.byref(prm), C.c_void_p(out.data_ptr()), ws.handle, stream)
        if rc:
            raise RuntimeError(L.icr_last_error().decode())

    # N > 1: the path's only exchange is that every rank receives all rewards: one NCCL all-gather per step.
    # ICR_BENCH_EXCHANGE=peer selects the fused variant instead (the reward kernel stores every reward straight into all
    # ranks' gathered vectors through peer memory and a step needs one device-side barrier); measured equal at N=2
    # (the all-gather already runs under the next step's first kernel), so the plain collective stays the default.
    peer = None
    if world > 1 and os.environ.get("ICR_BENCH_EXCHANGE", "nccl") == "peer":
        ok = torch.ones(1, device=dev)
        try:
            peer = sharding.PeerGather(V, dev)
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
    cabi = None
    if world > 1 and os.environ.get("ICR_BENCH_EXCHANGE") == "cabi":        # the same pipeline with the C ABI's own collective call
        cabi = sharding.NcclExchange(dev)
        exchange = "icr_exchange_rewards (C ABI: ncclAllGather + ncclAllReduce in one group) on a side stream under the next step's kernels"
    pipe = sharding.PipelinedGather(V, dev, gather=cabi) if world > 1 and peer is None and os.environ.get("ICR_BENCH_EXCHANGE") != "none" else None

    def drain():
        if pipe is not None:
            pipe.drain()

    def step():
        if pipe is not None:
            device_call(pipe.next_buffer())
            pipe.submit(ema.update)
            return
        if peer is not None:
            tg, fl, stp = peer.begin()
            rc = L.icr_compute_rewards_gather(p_i, p_l, 0, p_s, B, S, H, W, C.byref(prm), C.c_void_p(rewards.data_ptr()),
                                              C.c_void_p(tg.data_ptr()), C.c_void_p(fl.data_ptr()), C.c_ulonglong(stp), world, rank,
                                              ws.handle, stream)
            if rc:
                raise RuntimeError(L.icr_last_error().decode())
            reward.ema_update_gathered(ema, peer)
            return
        device_call(rewards)
        if world > 1 and os.environ.get("ICR_BENCH_EXCHANGE") != "none":      # "none": diagnosis only, prints an invalid line
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

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    # A batch that fits the 126 MB L2 (config 2: 51 MB) is timed step by step with the L2 flushed in between (a 512 MB
    # memset outside the timed pairs), as BASELINE.md section 3 asks; larger batches flush it by themselves.
    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev) if wl.step_bytes() < (200 << 20) else None

    def timed(fn, steps, after=lambda: None):
        sync()
        if flush_buf is not None:
            pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for a, b in pairs:
                flush_buf.fill_(1)
                a.record()
                fn()
                after()
                b.record()
            sync()
            return max_over_ranks(sum(a.elapsed_time(b) for a, b in pairs))
        e0.record()
        for _ in range(steps):
            fn()
        after()
        e1.record()
        sync()
        return max_over_ranks(e0.elapsed_time(e1))

    for _ in range(args.warmup):
        step()
    ws.check()
    sync()
    L.icr_launch_count(1)
    with ClockSampler(local) as clk:
        ms = timed(step, args.steps, drain)
    launches = int(L.icr_launch_count(1))
    value = wl.global_batch * args.steps / (ms * 1e-3)

    # ---- e2e: host buffers through the C ABI, H2D + D2H inside the timed region
    pi, pl, ps = torch.from_numpy(wl.images).pin_memory(), torch.from_numpy(wl.labels).pin_memory(), torch.from_numpy(wl.sizes).pin_memory()
    pr = torch.empty(V, dtype=torch.float64).pin_memory()

    def step_host():
        rc = L.icr_compute_rewards_host(C.c_void_p(pi.data_ptr()), C.c_void_p(pl.data_ptr()), 0, C.c_void_p(ps.data_ptr()), B, S, H, W,
                                        C.byref(prm), C.c_void_p(pr.data_ptr()), ws.handle, stream)
        if rc:
            raise RuntimeError(L.icr_last_error().decode())
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(3):
        step_host()
    e2e_ms = timed(step_host, e2e_steps)
    e2e_value = wl.global_batch * e2e_steps / (e2e_ms * 1e-3)
    device_call(rewards)                                             # untimed: the device entry's answer
    want = rewards.cpu().numpy()
    assert np.array_equal(pr.numpy(), want), "host-buffer entry and device entry disagree"

    # the same with what a data loader holds before its float conversion: u8 planes and (when the ids fit) u16 labels
    u8 = np.clip(np.rint(wl.images), 0, 255).astype(np.uint8)
    assert np.array_equal(u8.astype(np.float32), wl.images), "the synthetic planes hold whole bytes"
    small = wl.labels.min() >= 0 and wl.labels.max() < 65536
    pu = torch.from_numpy(u8).pin_memory()
    plu = torch.from_numpy(wl.labels.astype(np.uint16)).pin_memory() if small else pl
    del u8

    def step_host_u8():
        rc = L.icr_compute_rewards_host_u8(C.c_void_p(pu.data_ptr()), 0, C.c_void_p(plu.data_ptr()), 2 if small else 0, C.c_void_p(ps.data_ptr()),
                                           B, S, H, W, C.byref(prm), C.c_void_p(pr.data_ptr()), ws.handle, stream)
        if rc:
            raise RuntimeError(L.icr_last_error().decode())
    for _ in range(3):
        step_host_u8()
    u8_ms = timed(step_host_u8, e2e_steps)
    assert np.array_equal(pr.numpy(), want), "u8 host-buffer entry and device entry disagree"
    u8_bytes = int(pu.numel() + plu.numel() * plu.element_size() + wl.sizes.nbytes)

    # what the link alone takes for the same pinned buffers: e2e cannot go below this
    scratch_i, scratch_l = torch.empty_like(ti), torch.empty_like(tl)

    def copies():
        scratch_i.copy_(pi, non_blocking=True)
        scratch_l.copy_(pl, non_blocking=True)
    copies()
    copy_ms = timed(copies, 5) / 5
    del scratch_i, scratch_l, pu, plu, pi, pl

    # the reference-facing C++ call itself (libicr_host.so: compute_rewards_batched over libtorch tensors, inputs on the
    # device as src/fcn/training.cpp:174 has them; output allocation, the retry check and its stream sync inside)
    e2e_cpp = None
    host_lib = os.path.join(ROOT, "image-compression-cpp_b200", "libicr_host.so")
    if os.path.exists(host_lib):
        HL = C.CDLL(host_lib)
        HL.icrh_last_error.restype = C.c_char_p
        out_cpp = torch.empty(V, dtype=torch.float64, device=dev)

        def step_cpp():
            if S == 1:
                rc = HL.icrh_compute_rewards(p_i, 1, p_l, 0, p_s, B, H, W, local, 1, 4, C.c_float(0.012167), C.c_float(18.0), C.c_float(0.1),
                                             C.c_double(9.308622), 1, C.c_double(0.5), C.c_void_p(out_cpp.data_ptr()))
            else:
                rc = HL.icrh_compute_rewards_multi(p_i, p_l, 0, p_s, B, S, H, W, local, C.c_void_p(out_cpp.data_ptr()))
            if rc:
                raise RuntimeError(HL.icrh_last_error().decode())
        for _ in range(3):
            step_cpp()
        cpp_ms = timed(step_cpp, e2e_steps)
        assert np.array_equal(out_cpp.cpu().numpy(), want), "C++ host layer and C ABI disagree"
        e2e_cpp = {"value": wl.global_batch * e2e_steps / (cpp_ms * 1e-3), "unit": "images/sec", "ms_per_step": cpp_ms / e2e_steps,
                   "api": "libicr_host.so compute_rewards_batched (libtorch tensors on the device, as the reference's caller holds them; "
                          "allocation of the result, icr_check_detail and its stream sync inside the call)"}

    # ---- roofline of the dominant kernel: per-kernel device times from events between the launches
    L.icr_workspace_set_profile(ws.handle, 1)
    kms = np.zeros((args.steps, 4), np.float32)
    for i in range(args.steps):
        device_call(rewards)
        L.icr_profile_read(ws.handle, kms[i].ctypes.data_as(C.c_void_p))
    L.icr_workspace_set_profile(ws.handle, 0)
    kavg = kms.mean(0).astype(np.float64)
    names = ["prep_kernel", "narrow_kernel" if os.environ.get("ICR_ESTIMATOR") == "reg" else "estimate_kernel", "big_kernel", "reserved"]
    top = int(np.argmax(kavg))
    peak, peak_src = peaks()
    alg_bytes = wl.bytes_per_image() * B
    achieved = float(alg_bytes / (kavg[top] * 1e-3) / 1e9)
    path_gbs = float(alg_bytes / (ms / args.steps * 1e-3) / 1e9)

    traffic, traffic_src = None, None
    try:                          # dram bytes per launch of the same kernel from the committed ncu --set full capture of this command
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:
            tj = json.load(f)
        ent = tj.get(f"cfg{args.config}_B{B}")
        if ent:
            traffic, traffic_src = ent.get(names[top]), f"profiles/r02_traffic.json cfg{args.config}_B{B} ({tj.get('source', 'ncu --set full, per launch')})"
    except Exception:
        pass
    mb = wl.step_bytes() / 1e6
    line = {
        "metric": METRIC, "value": value, "unit": "images/sec", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": wl.name, "per_gpu_batch": B, "global_batch": wl.global_batch, "label_maps_per_image": S,
                   "mean_segments_per_label_map": wl.Kmean,
                   "l2": (f"inputs ({mb:.0f} MB/step/GPU) exceed L2; no flush" if flush_buf is None else
                          f"inputs ({mb:.0f} MB/step/GPU) fit L2: every step timed on its own with a 512 MB memset before it"),
                   "parallelism": f"dp{world}: images sharded, rewards exchanged by {exchange}, EMA on the gathered vector" if world > 1 else "single GPU"},
        "e2e": {"value": e2e_value, "unit": "images/sec", "h2d_bytes_per_step": wl.step_bytes(),
                "d2h_bytes_per_step": int(pr.numel() * 8), "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                "h2d_copy_alone_ms": copy_ms, "numa_node": numa,
                "api": "icr_compute_rewards_host (C ABI, pinned host buffers; the batch is copied in four pieces on a second stream "
                       "under the kernels of the piece before)"},
        "e2e_u8": {"value": wl.global_batch * e2e_steps / (u8_ms * 1e-3), "unit": "images/sec", "h2d_bytes_per_step": u8_bytes,
                   "d2h_bytes_per_step": int(pr.numel() * 8), "ms_per_step": u8_ms / e2e_steps,
                   "api": "icr_compute_rewards_host_u8 (u8 planes + " + ("u16" if small else "int32") + " labels over the link, widened on the "
                          "device; rewards bit-identical to e2e's)"},
        "gpu_launches": launches,
        "roofline": {"bound": "hbm", "kernel": names[top], "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": peak_src, "algorithmic_bytes_per_launch": float(alg_bytes),
                     "kernel_ms": {n: float(k) for n, k in zip(names, kavg)}, "whole_path_gbs": path_gbs, "whole_path_frac": path_gbs / peak,
                     "note": "the estimator is bound by the ALU pipe (packed-byte filter arithmetic, about 52 ALU-pipe instructions per "
                             "pixel), not by HBM: see DESIGN.md section 3.2 and profiles/r02_ncu.md"},
        "clocks": clk.summary(),
    }
    if e2e_cpp:
        line["e2e_cpp"] = e2e_cpp
    if world == 1 and args.config == 5:
        # Next to the headline, whose rectangles all fill their bboxes (the estimator then never reads a label): the same
        # batch with a third of each map's rectangles melted into a background that touches all four borders, which is what
        # multicut output looks like -- masked segments on the warp path, one whole-image Rule M segment per map on the
        # big-segment kernel.
        lab2 = wl.labels[:wl.pool].copy()
        for b in range(lab2.shape[0]):
            ids = np.unique(lab2[b])
            bg = np.random.default_rng(b).choice(ids, size=len(ids) // 3, replace=False)
            lab2[b][np.isin(lab2[b], bg)] = -5
            lab2[b][0, :], lab2[b][-1, :], lab2[b][:, 0], lab2[b][:, -1] = -5, -5, -5, -5
        reps = -(-B // wl.pool)
        tl2 = torch.from_numpy(np.ascontiguousarray(np.tile(lab2, (reps, 1, 1))[:B])).to(dev)
        r2 = torch.empty(V, dtype=torch.float64, device=dev)

        def step_bg():
            rc = L.icr_compute_rewards(p_i, C.c_void_p(tl2.data_ptr()), 0, p_s, B, S, H, W, C.byref(prm), C.c_void_p(r2.data_ptr()), ws.handle, stream)
            if rc:
                raise RuntimeError(L.icr_last_error().decode())
        for _ in range(3):
            step_bg()
        ws.check()
        bg_steps = max(3, min(args.steps, 10))
        bg_ms = timed(step_bg, bg_steps) / bg_steps
        line["irregular"] = {"workload": "the same batch with a third of each map's rectangles melted into a border-touching background segment",
                             "value": B / (bg_ms * 1e-3), "unit": "images/sec", "ms_per_step": bg_ms,
                             "mean_segments_per_label_map": float(np.mean([len(np.unique(m)) for m in lab2[:16]]))}
        del tl2, r2
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb, enc = cpu_baseline(wl)
        # per label map, the encoder's bytes next to the estimator's sum of S_k (a model of PNG size,
        # not an encoder: equality is not expected, correlation is)
        n = wl.pool
        device_call(rewards)
        torch.cuda.synchronize()
        K, _, _, seg = [t.cpu().numpy() for t in ws.tables()]
        est = np.array([seg[v, :K[v]].sum() for v in range(n * S)])
        cb["imencode_vs_estimate"] = {"label_maps": int(n * S), "pearson_r": float(np.corrcoef(enc, est)[0, 1]) if n * S > 1 else None,
                                      "sum_imencode_bytes": float(enc.sum()), "sum_estimated_bytes": float(est.sum()),
                                      "mean_ratio_estimate_over_imencode": float(np.mean(est / enc))}
        line["cpu_baseline"] = cb
    if rank == 0:
        print(json.dumps(line), file=args.out, flush=True)
    if world > 1:
        if cabi is not None:
            cabi.close()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
