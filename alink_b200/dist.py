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
