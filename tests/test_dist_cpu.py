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
