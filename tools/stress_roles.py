"""Race hunting for the warp-specialised tensor-core head: slow one role (or a set of roles) down per loop
iteration (ALINK_TC_DBG_DELAY="<role bits>:<ns>", csrc/head_tc.cuh) and require the utilities of every pair to be
bit-identical to the undisturbed run. A latent ordering bug shows up as a mismatch, a trap or a hang.
Roles: 1 weight producer, 2 MMA issuer, 8 epilogue, 16 gather producers, 32 peer relay, 128 L2 prefetcher."""
import os
import sys

import numpy as np

os.environ["ALINK_LIB_DEBUG"] = "1"  # the delay hooks only exist in the debug build of the library
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import _ffi, synth  # noqa: E402


def main():
    ctx = _ffi.Context(0)
    N = int(os.environ.get("STRESS_N", "120000"))
    pool = synth.make_pool(2, N=N, M=20000)
    ctx.set_table(0, pool.table)
    ctx.set_heads(0, pool.heads, compute=_ffi.BF16_TC)
    os.environ.pop("ALINK_TC_DBG_DELAY", None)
    per0, _ = ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
    bad = 0
    for ns in (2000, 20000):
        for bits in (1, 2, 8, 16, 32, 128, 1 | 8, 2 | 32, 8 | 32, 2 | 16, 1 | 2 | 8 | 16 | 32 | 128, 8 | 16, 1 | 32, 16 | 128):
            os.environ["ALINK_TC_DBG_DELAY"] = f"{bits}:{ns}"
            per, _ = ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
            same = np.array_equal(per.view(np.uint32), per0.view(np.uint32))
            bad += not same
            print(f"roles={bits:3d} delay={ns:6d} ns  identical={same}", flush=True)
    for bits, ns in ((128, 300000), (16, 200000), (8 | 128, 100000)):  # observers / producers several tiles behind
        os.environ["ALINK_TC_DBG_DELAY"] = f"{bits}:{ns}"
        per, _ = ctx.score_indexed(0, 0, 0, pool.ia, pool.ib, pool.N, pool.heads.K, pool.heads.C)
        same = np.array_equal(per.view(np.uint32), per0.view(np.uint32))
        bad += not same
        print(f"roles={bits:3d} delay={ns:6d} ns  identical={same}", flush=True)
    os.environ.pop("ALINK_TC_DBG_DELAY", None)
    print("STRESS", "FAILED" if bad else "OK")
    ctx.close()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
