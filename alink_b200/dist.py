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
