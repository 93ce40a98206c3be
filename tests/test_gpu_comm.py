"""The cross-shard exchange behind the C ABI (alink_comm_init / alink_query_sharded, csrc/comm.cuh) on REAL GPUs, with no
torch.distributed anywhere: two processes, one GPU each, the NCCL unique id shipped through the spawn arguments. Every rank
must return the single-GPU selection bit for bit (keys carry global indices). Needs >= 2 GPUs (skipped on the
one-GPU box; run with `gpurun --gpus 2`)."""
import multiprocessing as mp

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _worker(rank, world, uid, q, push):
    try:
        import os
        os.environ["ALINK_COMM_PUSH"] = str(push)  # read once, at the first alink_comm_init of the process
        from alink_b200 import _ffi, synth
        from alink_b200.dist import shard_range
        pool = synth.make_pool(2, N=300_001)  # ragged: the shards differ in size
        ctx = _ffi.Context(rank)
        ctx.set_table(0, pool.table)
        ctx.set_heads(0, pool.heads, compute=_ffi.FP32_REFINE)
        whole = ctx.query(0, 0, 0, pool.ia, pool.ib, pool.N, "uncertainty", 1000)
        ctx.comm_init(uid, rank, world)
        lo, hi = shard_range(pool.N, rank, world)
        out = {}
        for measure in ("uncertainty", "neg_margin"):
            ref = ctx.query(0, 0, 0, pool.ia, pool.ib, pool.N, measure, 1000)
            got = ctx.query_sharded(0, 0, 0, np.ascontiguousarray(pool.ia[lo:hi]), np.ascontiguousarray(pool.ib[lo:hi]), hi - lo,
                                    measure, 1000, base=lo)
            out[measure] = all(np.array_equal(a, b) for a, b in zip(ref, got))
        st = ctx.comm_push_state()  # (enabled, used): the one-launch peer push + merge of csrc/comm.cuh
        out["push_consistent"] = st == (bool(push), bool(push))
        print(f"rank {rank}: peer push enabled/used = {st}", flush=True)
        rep = ctx.refine_report()  # filled after the step's single synchronisation (the verdict travelled with the keys)
        out["report"] = bool(rep["version"] == 2 and rep["rounds"] == 1 and rep["n_sample"] >= 4096 and rep["n_exact"] >= 1000
                             and 0 < rep["err1_max"] <= rep["eps1_used"])
        # rank 0 verifies against a bound far too tight: its verdict is "not ok", EVERY rank must learn that from the gathered
        # verdicts, rank 0 repeats its query with widened bounds, and the exchange is repeated — same selection, no deadlock
        if rank == 0:
            ctx.set_refine_eps(2e-5)
        got = ctx.query_sharded(0, 0, 0, np.ascontiguousarray(pool.ia[lo:hi]), np.ascontiguousarray(pool.ib[lo:hi]), hi - lo,
                                "uncertainty", 1000, base=lo)
        rep = ctx.refine_report()
        ctx.set_refine_eps(0.0)
        out["one_rank_not_ok"] = bool(all(np.array_equal(a, b) for a, b in zip(whole, got))
                                      and (rank != 0 or push or rep["rounds"] > 1 or rep["version"] == 1))
        # a shard smaller than B: 600 pairs on rank 1, the rest on rank 0
        cut = pool.N - 600
        lo2, hi2 = (0, cut) if rank == 0 else (cut, pool.N)
        got = ctx.query_sharded(0, 0, 0, np.ascontiguousarray(pool.ia[lo2:hi2]), np.ascontiguousarray(pool.ib[lo2:hi2]), hi2 - lo2,
                                "uncertainty", 1000, base=lo2)
        out["small_shard"] = all(np.array_equal(a, b) for a, b in zip(whole, got))
        # the A-LINK rule with thresholds selected on the device across the ranks == the single-shot rule on the whole pool
        rng = np.random.default_rng(5)
        n, T = 40_001, 5
        p1 = rng.random(n).astype(np.float32)
        p2 = np.clip(p1[None, :] + rng.normal(0, 0.2, (T, n)), 0, 1).astype(np.float32)
        p2[:, 100:140] = p2[:, 60:100]          # ties across ...
        p1[100:140] = p1[60:100]
        y = (rng.random(n) < 0.5).astype(np.float32)
        for blind in (False, True):
            ridx, ract = ctx.disparity_select(p2, p1, 0.25, 0.05, blind, y)
            lo3, hi3 = shard_range(n, rank, world)
            gidx, gact = ctx.disparity_select_comm(np.ascontiguousarray(p2[:, lo3:hi3]), np.ascontiguousarray(p1[lo3:hi3]), lo3, n, 0.25,
                                                   0.05, blind, np.ascontiguousarray(y[lo3:hi3]))
            out[f"rule_blind{int(blind)}"] = bool(np.array_equal(gidx, ridx[(ridx >= lo3) & (ridx < hi3)]) and gact == ract)
        ctx.comm_destroy()
        ctx.close()
        q.put((rank, out, None))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, None, traceback.format_exc() + str(e)))


@pytest.mark.parametrize("push", [0, 1])  # the key exchange: ncclAllGather + merge (default) | one-launch peer push + merge
def test_two_ranks_merge_to_the_single_gpu_selection_without_torch_distributed(push):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from alink_b200 import _ffi
    uid = _ffi.Context.comm_unique_id()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    procs = [mpc.Process(target=_worker, args=(r, 2, uid, q, push)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, out, err in res:
        assert err is None, f"rank {rank}: {err}"
        assert out == {"uncertainty": True, "neg_margin": True, "push_consistent": True, "report": True, "one_rank_not_ok": True,
                       "small_shard": True, "rule_blind0": True, "rule_blind1": True}, (rank, out)
