"""Data-parallel plumbing of the reward path: contiguous image shards per rank, and the path's only exchange --
an all-gather of the per-image rewards (the reference is single-GPU,
include/ema_baseline.hpp:13 takes the mean of the whole batch).  Once every rank holds the whole reward vector in image
order, Rule E runs on it locally: one collective per step, and the EMA value is bit-identical to the single-GPU one
because the sum runs over the same values in the same order.  `exchange` keeps the all-reduce of {sum, count} for
callers that only want the statistics.
torch.distributed is the transport (NCCL over NVLink on the GPU box, gloo in the CPU tests); nothing here computes.
`PeerGather` removes even that collective on one NVLink box: every rank maps the others' gathered vectors
(torch symmetric memory) and the reward kernel stores each reward straight into all of them
(`icr_compute_rewards_gather`); a step then needs one device-side barrier instead of an all-gather."""
import torch
import torch.distributed as dist


def shard_range(B, world, rank):
    """Contiguous block of images owned by `rank`; the first B % world ranks take one extra image."""
    base, extra = divmod(B, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_rewards(rewards_local, B_total, group=None):
    """rewards_local f64 [b_r] -> rewards of the whole batch f64 [B_total] in image order; ONE all-gather.
    Ragged shards are padded to the largest shard."""
    world = dist.get_world_size(group)
    sizes = [shard_range(B_total, world, r)[1] - shard_range(B_total, world, r)[0] for r in range(world)]
    mx = max(sizes)
    pad = rewards_local
    if rewards_local.numel() < mx:
        pad = torch.zeros(mx, dtype=rewards_local.dtype, device=rewards_local.device)
        pad[:rewards_local.numel()] = rewards_local
    gathered = torch.empty(mx * world, dtype=rewards_local.dtype, device=rewards_local.device)
    dist.all_gather_into_tensor(gathered, pad.contiguous(), group=group)
    if all(s == mx for s in sizes):
        return gathered
    return torch.cat([gathered[r * mx:r * mx + sizes[r]] for r in range(world)])


def exchange(rewards_local, sum_count_local, B_total, group=None):
    """rewards_local f64 [b_r] (+ its {sum, count} f64 [2]) -> (rewards of the whole batch f64 [B_total] in image order,
    global {sum, count}).  Ragged shards are padded to the largest shard for the all-gather."""
    world = dist.get_world_size(group)
    sizes = [shard_range(B_total, world, r)[1] - shard_range(B_total, world, r)[0] for r in range(world)]
    mx = max(sizes)
    pad = rewards_local
    if rewards_local.numel() < mx:
        pad = torch.zeros(mx, dtype=rewards_local.dtype, device=rewards_local.device)
        pad[:rewards_local.numel()] = rewards_local
    gathered = torch.empty(mx * world, dtype=rewards_local.dtype, device=rewards_local.device)
    dist.all_gather_into_tensor(gathered, pad.contiguous(), group=group)
    sc = sum_count_local.clone()
    dist.all_reduce(sc, op=dist.ReduceOp.SUM, group=group)
    if all(s == mx for s in sizes):
        return gathered, sc
    return torch.cat([gathered[r * mx:r * mx + sizes[r]] for r in range(world)]), sc


class PipelinedGather:
    """The all-gather of step s and whatever consumes it (Rule E) run on a side stream while the caller's stream already
    computes step s+1; two reward buffers alternate.  Ranks then do not wait for each other inside every step -- measured
    on two B200s: 0.2936 ms per step with the exchange in line, 0.2815 ms pipelined, 0.2832 ms on one GPU alone.
    Per step: `buf = next_buffer()`, queue the kernels that fill `buf` on the current stream, `submit(consumer)`;
    `drain()` before reading results or closing a timed region."""

    def __init__(self, n_local, device, group=None, dtype=torch.float64, gather=None):
        self.group = group
        # gather(gathered, local) queues the exchange on the CURRENT stream; default: torch.distributed's all-gather
        self.gather = gather or (lambda g, l: dist.all_gather_into_tensor(g, l, group=self.group))
        self.world = dist.get_world_size(group)
        self.side = torch.cuda.Stream(device=device)
        self.local = [torch.empty(n_local, dtype=dtype, device=device) for _ in range(2)]
        self.gathered = [torch.empty(n_local * self.world, dtype=dtype, device=device) for _ in range(2)]
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.done = [torch.cuda.Event() for _ in range(2)]
        self.n = 0

    def next_buffer(self):
        par = self.n & 1
        if self.n >= 2:
            torch.cuda.current_stream().wait_event(self.done[par])      # the exchange that read this buffer two steps ago is over
        return self.local[par]

    def submit(self, consumer):
        par = self.n & 1
        self.ready[par].record(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.ready[par])
            self.gather(self.gathered[par], self.local[par])
            consumer(self.gathered[par])
            self.done[par].record(self.side)
        self.n += 1

    def drain(self):
        torch.cuda.current_stream().wait_stream(self.side)


class PeerGather:
    """Gathered reward vectors of all ranks, mapped into every rank (symmetric memory over NVLink), plus one flag per
    rank.  A step: `begin()` -> pass its three values to icr_compute_rewards_gather (the reward kernel stores every reward
    into all ranks' vectors and, when its last CTA is done, sets this rank's flag everywhere) -> `EMABaseline.
    update_gathered(peer)` waits on the device for all flags of the step, then reads the vector.  No collective, no
    barrier kernel.  Two vectors alternate by step, so a rank that runs ahead cannot overwrite what a slower rank still
    reads: it can only start step s+2 after it has seen every rank's flag of step s+1, which a rank raises after its
    reader of step s (earlier on the same stream) has finished.
    Raises if symmetric memory cannot be set up; callers fall back to `gather_rewards` (one NCCL all-gather)."""

    def __init__(self, n_local, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        self.world, self.rank, self.n_local = dist.get_world_size(group), dist.get_rank(group), int(n_local)
        self.n_total = self.n_local * self.world
        self.buf = symm_mem.empty(2 * self.n_total + self.world, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, group.group_name)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        i64 = dict(dtype=torch.int64, device=device)
        self._targets = [torch.tensor([p + par * self.n_total * 8 for p in ptrs], **i64) for par in (0, 1)]
        self.flag_ptrs = torch.tensor([p + 2 * self.n_total * 8 for p in ptrs], **i64)
        self.flags = self.buf[2 * self.n_total:].view(torch.int64)           # this rank's flags, one per rank
        self.step = 0
        torch.cuda.synchronize(device)
        self.handle.barrier(channel=0)                                       # every rank's zeros are in place
        torch.cuda.synchronize(device)

    def begin(self):
        """Start a step: (device array of the ranks' gathered vectors, device array of the ranks' flag arrays, step)."""
        self.step += 1
        return self._targets[self.step & 1], self.flag_ptrs, self.step

    def gathered(self):
        """This rank's gathered vector of the current step (valid once the flags of the step are up)."""
        par = self.step & 1
        return self.buf[par * self.n_total:(par + 1) * self.n_total]

    def barrier(self):
        """Device-side barrier across ranks on the current stream (for callers that pass no flags)."""
        self.handle.barrier(channel=0)


class NcclExchange:
    """The exchange through the C ABI itself (icr_exchange_rewards: all-gather of the rewards + all-reduce of {sum, count}
    in one NCCL group on the caller's stream), on a communicator of its own: rank 0 draws the id, torch.distributed (any
    backend) carries its 128 bytes to the others.  With world == 1 no process group is needed."""

    def __init__(self, device, group=None):
        import ctypes as C
        from . import reward
        self.C, self.L = C, reward.lib()
        self.device = torch.device(device)
        init = dist.is_available() and dist.is_initialized()
        self.world = dist.get_world_size(group) if init else 1
        self.rank = dist.get_rank(group) if init else 0
        ident = (C.c_ubyte * 128)()
        if self.rank == 0:
            self._check(self.L.icr_nccl_unique_id(ident))
        if self.world > 1:
            backend = dist.get_backend(group)
            t = torch.tensor(list(ident), dtype=torch.uint8, device=self.device if backend == "nccl" else "cpu")
            dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            ident = (C.c_ubyte * 128)(*t.cpu().tolist())
        self.comm = C.c_void_p()
        with torch.cuda.device(self.device):
            self._check(self.L.icr_nccl_comm_create(C.byref(self.comm), ident, self.world, self.rank))
        self.sum_count = torch.zeros(2, dtype=torch.float64, device=self.device)

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.icr_last_error().decode())

    def __call__(self, gathered, local):
        """Queue the exchange on the current stream: gathered f64 [world*n] <- every rank's local f64 [n]; self.sum_count
        receives the global {sum, count}."""
        C = self.C
        with torch.cuda.device(self.device):
            self._check(self.L.icr_exchange_rewards(self.comm, C.c_void_p(local.data_ptr()), C.c_int(local.numel()),
                                                    C.c_void_p(gathered.data_ptr()) if gathered is not None else None,
                                                    C.c_void_p(self.sum_count.data_ptr()),
                                                    C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def close(self):
        if self.comm:
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                self.L.icr_nccl_comm_destroy(self.comm)
            self.comm = self.C.c_void_p()
