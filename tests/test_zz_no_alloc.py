"""alink_topk_values_noalloc (include/alink_b200.h): the top-B selection with its scratch in a static device array, for
callers that cannot allocate. Through the C ABI, against the existing alink_topk_values on the same inputs: the result is
integer / bit work, so the bar is equality of keys, indices and utility bits — on two calls in a row with DIFFERENT inputs
(the array keeps the first call's state: a stale threshold or slot counter would show in the second).
Collected last on purpose (a failure here must not hide other tests)."""
import numpy as np
import pytest

from alink_b200 import _ffi

pytestmark = pytest.mark.gpu


def _values(seed, N):
    rng = np.random.default_rng(seed)
    v = rng.random(N, dtype=np.float32)
    v[rng.integers(0, N, N // 8)] = np.float32(0.75)  # a plateau of ties: the lowest pair index must win
    v[rng.integers(0, N, 64)] = np.float32(1.0)
    return v


def _run(ctx, v, B, base=0):
    import torch
    dv = torch.from_numpy(v).cuda()
    idx = torch.full((B,), -1, dtype=torch.int64, device="cuda")
    util = torch.full((B,), -1.0, dtype=torch.float32, device="cuda")
    keys = torch.zeros((B,), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    ctx.topk_values_noalloc(dv, B, idx, util, keys, base=base)
    ctx.sync()
    torch.cuda.synchronize()
    return idx.cpu().numpy(), util.cpu().numpy(), keys.cpu().numpy().view(np.uint64)


@pytest.mark.parametrize("N,B", [(300_000, 1000), (2_000_000, 4096), (5_000, 37), (40_000, 1)])
def test_matches_alink_topk_values_twice_in_a_row(ctx, N, B):
    for call, seed in enumerate((11, 12)):
        v = _values(seed + N, N)
        want = ctx.topk_values(v, B, base=7 * call)
        got = _run(ctx, v, B, base=7 * call)
        np.testing.assert_array_equal(got[2], want[2], err_msg=f"keys, call {call}")
        np.testing.assert_array_equal(got[0], want[0], err_msg=f"indices, call {call}")
        np.testing.assert_array_equal(got[1].view(np.uint32), want[1].view(np.uint32), err_msg=f"utilities, call {call}")
        # and against numpy: utility descending, index ascending
        order = np.lexsort((np.arange(N), -v.astype(np.float64)))[:B]
        np.testing.assert_array_equal(got[0] - 7 * call, order)


def test_outputs_are_optional_and_nothing_is_allocated(ctx):
    import torch
    v = _values(5, 500_000)
    dv = torch.from_numpy(v).cuda()
    keys = torch.zeros((256,), dtype=torch.int64, device="cuda")
    ctx.topk_values(v, 256)  # (the existing entry point grows the context's buffers; do it before the measurement)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        ctx.topk_values_noalloc(dv, 256, None, None, keys)
    ctx.sync()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "device memory changed across the no-allocation calls"
    np.testing.assert_array_equal(keys.cpu().numpy().view(np.uint64), ctx.topk_values(v, 256)[2])


def test_sizes_beyond_the_static_scratch_are_refused(ctx):
    import torch
    dv = torch.zeros((10_000,), dtype=torch.float32, device="cuda")
    keys = torch.zeros((8192,), dtype=torch.int64, device="cuda")
    with pytest.raises(_ffi.AlinkError, match="exceeds the static scratch"):
        ctx.topk_values_noalloc(dv, 4097, None, None, keys)
    ctx.topk_values_noalloc(dv, 4096, None, None, keys)  # the limit itself is taken; all-equal utilities -> indices 0..4095
    ctx.sync()
    k = keys.cpu().numpy().view(np.uint64)[:4096]
    np.testing.assert_array_equal(k, ctx.topk_values(np.zeros(10_000, np.float32), 4096)[2])
