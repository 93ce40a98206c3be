"""Where the N > 1 step goes: per rank, the wall time of the local alink_query (no exchange), of alink_query_sharded (exchange
+ merge), and of the local query with the ranks held in lock step by a barrier — run under torchrun (N = 2 / 8)."""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_stream(torch.cuda.Stream())
N, B, steps = 1_000_000, 1000, int(os.environ.get("PROBE_STEPS", 30))
pool = synth.make_pool(2, N=N)
ctx = _ffi.Context(int(os.environ["LOCAL_RANK"]))
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
uid = [_ffi.Context.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
ctx.comm_init(uid[0], rank, world)
tab = torch.from_numpy(pool.table).to(dev)
ia, ib = torch.from_numpy(pool.ia).to(dev), torch.from_numpy(pool.ib).to(dev)
ctx.set_table(0, tab)
ctx.set_heads(0, pool.heads, compute=_ffi.FP32_REFINE)
keys_dev = torch.zeros((B,), dtype=torch.int64, device=dev)


def timed(fn):
    for _ in range(5):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps * 1e3


local = timed(lambda: ctx.query(0, 0, 0, ia, ib, N, "uncertainty", B, base=rank * N))
shard = timed(lambda: ctx.query_sharded(0, 0, 0, ia, ib, N, "uncertainty", B, base=rank * N))
res = [None] * world
dist.all_gather_object(res, (rank, round(local, 4), round(shard, 4), ctx.comm_push_state()))
if rank == 0:
    print("rank, local query ms, sharded query ms, push (enabled, used):")
    for r in res:
        print("  ", r)
ctx.comm_destroy()
ctx.close()
dist.destroy_process_group()
