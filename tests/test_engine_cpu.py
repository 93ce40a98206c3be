"""Host-side bookkeeping of alink_b200/engine.py that needs no GPU (cache keys, table-slot ownership)."""
import gc

import numpy as np
import pytest

from alink_b200 import _ffi, engine


def test_group_cache_keys_are_never_reused():
    """Short-lived HeadSpecs (models built in a loop) must not collide: CPython reuses id() of freed objects."""
    keys = set()
    for seed in range(64):
        s = engine.HeadSpec(512, 512, 64, 2, "softmax")
        s.set(*[np.zeros(shp, np.float32) for shp in ((512, 512), (512,), (512, 64), (64,), (64, 2), (2,))])
        keys.add(engine._GroupCache.key_of([s]))
        del s
        gc.collect()
    assert len(keys) == 64


def test_key_changes_with_weights_and_mode():
    s = engine.HeadSpec(64, 128, 32, 2, "softmax")
    z = [np.zeros(shp, np.float32) for shp in ((64, 128), (128,), (128, 32), (32,), (32, 2), (2,))]
    s.set(*z)
    k1 = engine._GroupCache.key_of([s])
    s.set(*z)
    assert engine._GroupCache.key_of([s]) != k1


class _FakeCtx:
    def set_table(self, slot, emb):
        self.last = slot


def test_pool_slots_are_owned_until_released():
    ctx = _FakeCtx()
    t = np.zeros((4, 8), np.float32)
    ia = np.zeros((2,), np.int32)
    pools = [engine.PairPool(t, ia, ia, ctx=ctx) for _ in range(engine.PairPool.POOL_SLOTS)]
    assert sorted(p.slot_a for p in pools) == list(range(engine.PairPool.POOL_SLOTS))
    with pytest.raises(_ffi.AlinkError):  # a further pool would overwrite a live pool's table
        engine.PairPool(t, ia, ia, ctx=ctx)
    freed = pools[2].slot_a
    pools[2].close()
    p = engine.PairPool(t, ia, ia, ctx=ctx)
    assert p.slot_a == freed
    del pools[0]
    gc.collect()
    with pytest.raises(_ffi.AlinkError):  # two tables need two slots: only one is free
        engine.PairPool(t, ia, ia, table_b=t, ctx=ctx)
    del pools[0]
    gc.collect()
    q = engine.PairPool(t, ia, ia, table_b=t, ctx=ctx)
    assert q.slot_a != q.slot_b
