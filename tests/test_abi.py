"""The C-ABI library loads and exports every symbol include/alink_b200.h declares (no compute here)."""
import ctypes
import os
import re

import pytest

from alink_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "alink_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(alink_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _ffi.load()
    names = _declared_symbols()
    assert len(names) >= 19
    for n in names:
        assert hasattr(lib, n), f"libalink_b200.so does not export {n}"
    assert sorted(_ffi.SYMBOLS) == names, "alink_b200/_ffi.py binds a different symbol set than the header declares"


def test_abi_version():
    assert _ffi.load().alink_abi_version() == 1


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(_ffi.AlinkError, match="no CPU fallback"):
        _ffi.Context(0)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under alink_b200/ may import, link or execute it."""
    pkg = os.path.join(ROOT, "alink_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", "Makefile")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle_ffi" not in txt and "libalink_oracle" not in txt and "ref_numpy" not in txt, f
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f


def test_sass_is_sm100a():
    so = _ffi.SO_PATH
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_hot_kernels_are_tcgen05_native():
    """SASS of the built library (B200_PROFILING.md: tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, cp.async.bulk -> UBLKCP,
    tcgen05.commit -> UTCBAR, mbarrier -> SYNCS): every variant of the three tensor-core head kernels carries them, and
    none of them falls back to the legacy mma.sync path (HMMA). profiles/r2_sass_opcodes.txt is the committed histogram."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("sass_opcodes", os.path.join(ROOT, "tools", "sass_opcodes.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    per = mod.histogram(_ffi.SO_PATH)
    heads = {n: c for n, c in per.items() if "head_kernel" in n and "train_head" not in n}
    assert len(heads) >= 12 and any("split" in n for n in heads) and any("tcko" in n for n in heads)
    for name, cnt in heads.items():
        ops = "".join(cnt)
        assert any(k.startswith("UTCHMMA") for k in cnt), f"{name}: no tcgen05.mma (UTCHMMA) in SASS: {dict(cnt)}"
        assert any(k.startswith("LDTM") for k in cnt), f"{name}: no tcgen05.ld (LDTM)"
        assert any(k.startswith("UBLKCP") for k in cnt), f"{name}: no bulk copy (UBLKCP)"
        assert any(k.startswith("UTCBAR") for k in cnt) and cnt.get("SYNCS", 0) > 0, f"{name}: no tcgen05.commit / mbarrier"
        assert "HMMA" not in ops.replace("UTCHMMA", ""), f"{name}: legacy mma.sync path present"
    # the default filter of FP32_REFINE / BF16_TC: CTA pairs (cta_group::2) with multicast commits
    pair = [c for n, c in heads.items() if "ILb0ELi2ELb1ELb1E" in n]
    assert pair and any(k.startswith("UTCHMMA.2CTA") for k in pair[0]) and any("MULTICAST" in k for k in pair[0])
