"""N > 1 host logic on CPU: two gloo ranks shard the pool by pair index, each computes its shard's top-B
keys (with the CPU oracle standing in for the GPU step — this test is about dist.py, not the kernels),
all-gather + merge, and every rank must end with the single-shot global selection."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, N, B, out_dir):
    import torch.distributed as dist
    from alink_b200 import synth
    from alink_b200.dist import decode_keys, merge_shards, shard_range
    from oracle import oracle_ffi

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pool = synth.make_pool(1, N=N)
    lo, hi = shard_range(N, rank, world)
    ref = oracle_ffi.score(pool.table, pool.table, pool.ia[lo:hi], pool.ib[lo:hi], pool.heads, measure="vote_entropy",
                           threads=2)
    b_local = min(B, hi - lo)
    _, _, keys = oracle_ffi.topk(ref["util"], b_local, base=lo)  # global indices via base
    merged = merge_shards(keys, B, oracle_ffi.topk_keys, dist=dist)
    idx, util = decode_keys(merged)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), keys=merged, idx=idx, util=util)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("N,B", [(1001, 100), (300, 250)])
def test_two_rank_merge_equals_single_shot(tmp_path, N, B):
    import torch.multiprocessing as mp
    from alink_b200 import synth
    from oracle import oracle_ffi

    oracle_ffi.build()
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, N, B, str(tmp_path)), nprocs=world, join=True)
    pool = synth.make_pool(1, N=N)
    ref = oracle_ffi.score(pool.table, pool.table, pool.ia, pool.ib, pool.heads, measure="vote_entropy")  # heavy ties
    ridx, rutil, rkeys = oracle_ffi.topk(ref["util"], B)
    for r in range(world):
        z = np.load(tmp_path / f"r{r}.npz")
        np.testing.assert_array_equal(z["keys"], rkeys)
        np.testing.assert_array_equal(z["idx"], ridx)
        np.testing.assert_array_equal(z["util"], rutil)


def test_shard_ranges_partition_the_pool():
    from alink_b200.dist import shard_range
    for N in (0, 1, 7, 1000, 10_000_001):
        for world in (1, 2, 3, 8):
            edges = [shard_range(N, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == N
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


def test_decode_keys_roundtrip():
    from alink_b200.dist import decode_keys
    from oracle import oracle_ffi
    rng = np.random.default_rng(0)
    u = rng.standard_normal(64).astype(np.float32)
    u[:3] = [0.0, -0.0, 1.5]
    idx = rng.integers(0, 2**32 - 2, size=64)
    keys = np.array([oracle_ffi.pack_key(float(a), int(b)) for a, b in zip(u, idx)], dtype=np.uint64)
    di, du = decode_keys(keys)
    np.testing.assert_array_equal(di, idx)
    np.testing.assert_array_equal(du, np.where(u == 0, np.float32(0.0), u))


# ---- sharded A-LINK disparity rule (dist.sharded_disparity_select) ---------------------------------------
def _pack_keys(d, idx):
    """numpy twin of alink_pack_key (include/alink_math.h) for d >= 0."""
    bits = np.ascontiguousarray(d, np.float32).view(np.uint32).astype(np.uint64)
    o = bits | np.uint64(0x80000000)
    return (o << np.uint64(32)) | (np.uint64(0xFFFFFFFF) - idx.astype(np.uint64))


class _NumpyShard:
    """Stand-in for the two Context calls dist.sharded_disparity_select makes (the kernels are tested on the GPU;
    this file is about the host-side exchange)."""

    def disparity_hist(self, q, p1, base, prefix, pass_):
        from alink_b200.dist import RADIX_BITS, RADIX_SHIFTS
        shift, bits = RADIX_SHIFTS[pass_], RADIX_BITS[pass_]
        keys = _pack_keys(np.abs(q - p1), base + np.arange(len(q)))
        if pass_ > 0:
            keys = keys[(keys >> np.uint64(shift + bits)) == np.uint64(prefix)]
        return np.bincount(((keys >> np.uint64(shift)) & np.uint64((1 << bits) - 1)).astype(np.int64), minlength=2048)

    def disparity_select_sharded(self, p2, p1, base, thr, eps, blind, y):
        T, n = p2.shape
        cnt = np.zeros(n, np.int64)
        for t in range(T):
            if blind:
                cnt += (p2[t] >= 0.5) != (p1 >= 0.5)
            elif thr[t]:
                cnt += _pack_keys(np.abs(p2[t] - p1), base + np.arange(n)) >= thr[t]
        p = p1.astype(np.float64)
        q = (cnt == T) & ((p <= 0.5 - eps) | (p >= 0.5 + eps))
        sel = q if y is None else q & ((p1 >= 0.5) == (y >= 0.5))
        return base + np.nonzero(sel)[0].astype(np.int64), int(q.sum())


def _disp_inputs(N, T, seed=3):
    rng = np.random.default_rng(seed)
    p1 = rng.random(N).astype(np.float32)
    p2 = np.clip(p1[None, :] + 0.2 * rng.standard_normal((T, N)), 0, 1).astype(np.float32)
    p2[:, ::7] = p1[::7]  # exact ties at d = 0
    p2[0, 5:40] = p1[5:40] + np.float32(0.25)  # a plateau of equal disparities across the shard boundary
    y = (rng.random(N) < 0.5).astype(np.float32)
    return p2, p1, y


def _disp_worker(rank, world, port, N, T, ratio, blind, out_dir):
    import torch.distributed as dist
    from alink_b200.dist import shard_range, sharded_disparity_select
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p2, p1, y = _disp_inputs(N, T)
    lo, hi = shard_range(N, rank, world)
    idx, active = sharded_disparity_select(_NumpyShard(), np.ascontiguousarray(p2[:, lo:hi]), p1[lo:hi], lo, N, ratio,
                                           0.05, blind, y[lo:hi], dist=dist)
    np.savez(os.path.join(out_dir, f"d{rank}.npz"), idx=idx, active=active)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("N,T,ratio,blind", [(1003, 3, 0.25, False), (400, 2, 0.9, False), (257, 2, 0.25, True),
                                             (64, 2, 0.001, False)])
def test_two_rank_disparity_equals_single_shot(tmp_path, N, T, ratio, blind):
    import torch.multiprocessing as mp
    from oracle import oracle_ffi

    oracle_ffi.build()
    world, port = 2, _free_port()
    mp.spawn(_disp_worker, args=(world, port, N, T, ratio, blind, str(tmp_path)), nprocs=world, join=True)
    p2, p1, y = _disp_inputs(N, T)
    ridx, ractive = oracle_ffi.disparity_select(p2, p1, ratio, 0.05, blind, y)
    for r in range(world):
        z = np.load(tmp_path / f"d{r}.npz")
        np.testing.assert_array_equal(z["idx"], ridx)
        assert int(z["active"]) == ractive
