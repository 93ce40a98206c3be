"""dist.py — data-parallel query selection: shard the pool by pair index, merge per-shard top-B.

The reference is single-process; this is the build's multi-GPU layer.
One process per GPU (torchrun). Pairs are independent and top-B is decomposable
(global top-B is a subset of the union of per-shard top-B), so the ONLY exchange is one all-gather of B packed
u64 keys per rank (8 KB at B = 1000) followed by the same deterministic merge on every rank
(`alink_topk_keys`). Keys carry GLOBAL pair indices (base = shard offset), so the merged selection
is identical to the single-GPU one, ties included.
"""
from __future__ import annotations

import numpy as np


def shard_range(N: int, rank: int, world: int):
    """Contiguous pair-index range [lo, hi) of `rank`."""
    return (N * rank) // world, (N * (rank + 1)) // world


def merge_shards(local_keys: np.ndarray, B: int, merge_fn, dist=None, device=None):
    """All-gather each rank's candidate keys (uint64, sorted desc, <= B of them) and merge to the
    global top-B. merge_fn(all_keys uint64[n], B) -> uint64[B] runs the merge (GPU: Context.topk_keys).
    Without an initialised process group this is the identity merge of one shard."""
    import torch
    if isinstance(local_keys, torch.Tensor) and local_keys.is_cuda:
        # device path: the keys never visit the host before the merge. local_keys = int64 tensor of B slots
        # (bit patterns of the u64 keys, zero-padded); one NCCL all-gather, one merge, one 8*B-byte read-back.
        if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
            return local_keys[:B].cpu().numpy().view(np.uint64)
        world = dist.get_world_size()
        gathered = torch.empty((world * local_keys.numel(),), dtype=local_keys.dtype, device=local_keys.device)
        dist.all_gather_into_tensor(gathered, local_keys)
        return merge_fn(gathered, B)  # zero padding sorts below every real key
    local_keys = np.ascontiguousarray(local_keys, dtype=np.uint64)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_keys[:B]  # one shard: its sorted top-B is the answer
    world = dist.get_world_size()
    padded = np.zeros((B,), dtype=np.uint64)  # key 0 sorts below every real key
    padded[: len(local_keys)] = local_keys
    t = torch.from_numpy(padded.view(np.int64))
    if device is not None:
        t = t.to(device)
    gathered = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    allk = torch.stack(gathered).reshape(-1).cpu().numpy().view(np.uint64)
    allk = allk[allk != 0]
    return merge_fn(np.ascontiguousarray(allk), min(B, len(allk)))


class ShardedTableFeed:
    """Replicated embedding table fed by 1/world-sized PCIe uploads + one NVLink all-gather (SURVEY.md 8e, e2e path).

    Every rank needs the WHOLE table (pairs reference arbitrary rows), but uploading it over each rank's own PCIe link
    multiplies host-memory traffic by the world size (8 x 410 MB per step on one NUMA node: 0.42 e2e scaling efficiency in
    round 1). Here rank r uploads only rows [r M / W, (r + 1) M / W) from its pinned host copy into its slice of a device
    buffer, and one in-place `all_gather_into_tensor` over NCCL / NVSwitch completes the replica on every GPU. Both run
    on a side stream, so with two buffers the feed of step i + 1 overlaps the query of step i; `wait(buf)` makes the
    compute stream wait on the device (no host synchronisation). world == 1 degenerates to the plain upload.
    M is padded to a multiple of the world size (pad rows are never referenced)."""

    def __init__(self, ctx, M, D, dist=None, device=None, dtype=None, n_buffers=2, gather="nccl"):
        """gather: "nccl" = in-place ncclAllGather (a kernel: it needs SMs, which the persistent head kernel holds, so it
        runs after the filter and delays the split kernel's CTA pairs); "copy" = the buffers live in symmetric memory
        (torch.distributed._symmetric_memory: cuMem allocations mapped into every rank) and every rank PULLS the other
        ranks' slices with peer-to-peer copies on the copy engines over NVLink — no SMs, two device-side barriers per feed;
        "auto" = "copy" when symmetric memory can be set up, else "nccl". Measured on 8 B200s (bench.py e2e): 3.59 ms
        ("copy") against 3.54 ms ("nccl") per step — the refine stage no longer waits for SMs (0.50 instead of 0.70 ms), but
        the barrier kernels cannot start before the persistent filter kernel ends either, so the table is ready just as
        late; hence "nccl", the plainer path, is the default."""
        import torch
        self.ctx, self.dist, self.torch = ctx, dist, torch
        self.world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self.M, self.D = int(M), int(D)
        self.rows = (self.M + self.world - 1) // self.world  # rows per rank
        self.device = device if device is not None else torch.device("cuda", ctx.device)
        dtype = dtype or torch.float32
        self.mode, self.hdl = "nccl", [None] * n_buffers
        self.bufs = None
        if self.world > 1 and gather in ("auto", "copy"):
            try:
                import torch.distributed._symmetric_memory as symm
                bufs = [symm.empty((self.rows * self.world, self.D), dtype=dtype, device=self.device) for _ in range(n_buffers)]
                self.hdl = [symm.rendezvous(b, dist.group.WORLD) for b in bufs]
                if any(h is None for h in self.hdl):
                    raise RuntimeError("rendezvous returned no handle")
                self.bufs, self.mode = bufs, "copy"
            except Exception as e:  # noqa: BLE001 — any failure of the experimental API falls back to NCCL
                if gather == "copy":
                    raise
                self.why_nccl = repr(e)[:200]
        if self.bufs is None:
            self.bufs = [torch.empty((self.rows * self.world, self.D), dtype=dtype, device=self.device) for _ in range(n_buffers)]
        self.side = torch.cuda.Stream(device=self.device)
        self.ready = [torch.cuda.Event() for _ in range(n_buffers)]
        self.free = [torch.cuda.Event() for _ in range(n_buffers)]
        self.idx = [None] * n_buffers
        self.nidx = [0] * n_buffers
        self.bytes_h2d = self.rows * self.D * self.bufs[0].element_size()
        self.bytes_gather = self.bytes_h2d * (self.world - 1)  # received per rank over NVLink

    def shard_of(self, host_table):
        """This rank's rows of a host table (a view; pin it once with .pin_memory() for full PCIe rate)."""
        lo = self.rank * self.rows
        return host_table[lo:min(lo + self.rows, self.M)]

    def feed_async(self, buf, host_shard, host_indices=None):
        """Start filling buffer `buf` from this rank's pinned shard; returns at once. host_indices: optional pinned
        (ia, ib) int32 tensors of the SAME step — they ride on the side stream too (ahead of the table, so a query
        never waits for them behind the upload) and `wait` hands back their device copies."""
        torch = self.torch
        cur = torch.cuda.current_stream(self.device)
        if cur.cuda_stream == 0:
            raise RuntimeError("ShardedTableFeed needs a non-default compute stream (torch.cuda.set_stream(torch.cuda.Stream())): "
                               "the legacy default stream synchronises with the feed's side stream and serialises the overlap")
        self.free[buf].record(cur)           # queries already queued on the compute stream may still read the buffer
        full = self.bufs[buf]
        lo = self.rank * self.rows
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.free[buf])
            if host_indices is not None:
                if self.idx[buf] is None or self.idx[buf][0].numel() < host_indices[0].numel():
                    self.idx[buf] = tuple(torch.empty(t.numel(), dtype=torch.int32, device=self.device) for t in host_indices)
                for d, h_ in zip(self.idx[buf], host_indices):
                    d[:h_.numel()].copy_(h_, non_blocking=True)
                self.nidx[buf] = int(host_indices[0].numel())
            mine = full[lo:lo + self.rows]
            mine[:host_shard.shape[0]].copy_(host_shard, non_blocking=True)
            if self.world > 1 and self.mode == "copy":
                h = self.hdl[buf]
                h.barrier()                      # every rank's slice has landed in its own buffer
                for step in range(1, self.world):
                    r = (self.rank - step) % self.world
                    peer = h.get_buffer(r, full.shape, full.dtype)
                    full[r * self.rows:(r + 1) * self.rows].copy_(peer[r * self.rows:(r + 1) * self.rows], non_blocking=True)
                h.barrier()                      # nobody still reads a slice that the feed after next will overwrite
            elif self.world > 1:
                self.dist.all_gather_into_tensor(full, mine)  # in place: rank r's slice is already where it belongs
            self.ready[buf].record(self.side)

    def wait(self, buf, slot):
        """Compute stream waits (on the device) for the feed of `buf`, then the context's `slot` borrows it."""
        torch = self.torch
        torch.cuda.current_stream(self.device).wait_event(self.ready[buf])
        self.ctx.set_table(slot, self.bufs[buf][:self.M])
        if self.idx[buf] is not None:
            return tuple(t[:self.nidx[buf]] for t in self.idx[buf])
        return None


def decode_keys(keys: np.ndarray):
    """(pair index int64, utility fp32) from packed keys (include/alink_math.h: alink_pack_key)."""
    keys = np.asarray(keys, dtype=np.uint64)
    idx = (np.uint64(0xFFFFFFFF) - (keys & np.uint64(0xFFFFFFFF))).astype(np.int64)
    o = (keys >> np.uint64(32)).astype(np.uint32)
    neg = (o & np.uint32(0x80000000)) == 0
    bits = np.where(neg, ~o, o & np.uint32(0x7FFFFFFF)).astype(np.uint32)
    return idx, bits.view(np.float32)


def sharded_query(ctx, group, slot_a, slot_b, ia_local, ib_local, n_local, lo, measure, B, dist=None,
                  device=None):
    """Per-rank fused query on its shard (global indices via base = lo), then the key merge."""
    b_local = min(B, n_local)
    _, _, keys = ctx.query(group, slot_a, slot_b, ia_local, ib_local, n_local, measure, b_local, base=lo)
    merged = merge_shards(keys, B, ctx.topk_keys, dist=dist, device=device)
    idx, util = decode_keys(merged)
    return idx, util, merged


# ---- A-LINK disparity rule on a sharded pool (code/ALINK.py:170-201) -----------------------
RADIX_SHIFTS = (53, 42, 32, 21, 10, 0)  # csrc/topk.cuh: 11/11/10 bits of the value, 11/11/10 bits of the index part
RADIX_BITS = (11, 11, 10, 11, 11, 10)


def _all_reduce_sum(arr: np.ndarray, dist, device):
    """Sum a small int64 numpy array over the ranks (identity without a process group)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return arr
    import torch
    t = torch.from_numpy(arr.copy())
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def global_threshold_key(hist_fn, m: int, dist=None, device=None) -> int:
    """The m-th largest key (1-based) over ALL shards, by six radix passes.

    hist_fn(prefix, pass_) -> 2048 counters of this shard's keys that match `prefix` on the bits above the
    pass (Context.disparity_hist). One all-reduce of 16 KB per pass; every rank walks the summed histogram
    from the top bin down exactly like radix_pass_kernel's last CTA, so all ranks end with the same key."""
    prefix, remaining = 0, int(m)
    for p, (shift, bits) in enumerate(zip(RADIX_SHIFTS, RADIX_BITS)):
        local = np.asarray(hist_fn(prefix, p), dtype=np.int64)
        tot = _all_reduce_sum(local, dist, device)
        nb = 1 << bits
        cum = 0
        chosen = None
        for b in range(nb - 1, -1, -1):
            c = int(tot[b])
            if cum + c >= remaining:
                chosen = b
                break
            cum += c
        if chosen is None:
            raise ValueError("rank %d exceeds the number of keys" % m)
        prefix = chosen if p == 0 else ((prefix << bits) | chosen)
        remaining -= cum
    return prefix


def sharded_disparity_select(ctx, p2_local, p1_local, lo: int, n_total: int, disparity_ratio=0.25, eps=0.05,
                             blind_strategy=False, y_local=None, dist=None, device=None, gather=True):
    """A-LINK's selector when the pool is sharded by pair index: p2_local (T, n) / p1_local (n,) are this
    rank's pairs [lo, lo + n). Per noise type the threshold of argsort(d)[:int(n_total*ratio)] is found with
    global_threshold_key (no score leaves its GPU); membership, intersection over t, grey zone and oracle
    agreement stay local. Returns (queryIndices, ACTIVE_COUNT): global and identical on every rank when
    gather=True (one all-gather of the selected indices), else this shard's part and the global count."""
    T = int(p2_local.shape[0])
    m = int(n_total * disparity_ratio)  # int(len * ratio), code/ALINK.py:183
    thr = np.zeros((T,), np.uint64)
    if not blind_strategy and m > 0:
        for t in range(T):
            q = p2_local[t]
            thr[t] = np.uint64(global_threshold_key(lambda pre, ps: ctx.disparity_hist(q, p1_local, lo, pre, ps), m,
                                                    dist=dist, device=device))
    idx, active = ctx.disparity_select_sharded(p2_local, p1_local, lo, thr, eps, blind_strategy, y_local)
    active = int(_all_reduce_sum(np.array([active], np.int64), dist, device)[0])
    if not gather or dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return idx, active
    import torch
    world = dist.get_world_size()
    counts = np.zeros((world,), np.int64)
    counts[dist.get_rank()] = len(idx)
    counts = _all_reduce_sum(counts, dist, device)
    cap = int(counts.max()) if world else 0
    buf = np.full((max(cap, 1),), -1, np.int64)
    buf[: len(idx)] = idx
    t = torch.from_numpy(buf)
    if device is not None:
        t = t.to(device)
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    out = np.concatenate([parts[r].cpu().numpy()[: int(counts[r])] for r in range(world)])
    return out, active  # shards are contiguous index ranges in rank order: already ascending
