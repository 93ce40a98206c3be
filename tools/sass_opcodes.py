"""Per-kernel histogram of the Blackwell-native SASS opcodes in libalink_b200.so (cuobjdump -sass):
UTCHMMA* (tcgen05.mma), LDTM* (tcgen05.ld), UBLKCP* / UBLKPF* (cp.async.bulk / prefetch), UTCBAR* (tcgen05.commit),
SYNCS* (mbarrier), plus FFMA / HMMA counts for contrast.   python tools/sass_opcodes.py > profiles/r2_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "alink_b200", "csrc", "libalink_b200.so")
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UBLKCP[.\w]*|UBLKPF[.\w]*|UTCBAR[.\w]*|SYNCS[.\w]*|UTMALDG[.\w]*|HMMA[.\w]*|FFMA\b|UTCATOMSWS[.\w]*)")


def histogram(so=SO):
    out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
    per = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = per.setdefault(m.group(1), collections.Counter())
            continue
        if cur is None:
            continue
        m = PAT.search(line)
        if m:
            op = m.group(1)
            cur[op.split(".")[0] if op.startswith(("SYNCS", "FFMA")) else op] += 1
    return per


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    except Exception:
        return list(names)


if __name__ == "__main__":
    per = histogram()
    names = demangle(list(per))
    print(f"# {os.path.relpath(SO, ROOT)}: SASS opcode histogram per kernel (cuobjdump -sass); tcgen05.mma = UTC*MMA, tcgen05.ld = LDTM,")
    print("# cp.async.bulk = UBLKCP, tcgen05.commit = UTCBAR, mbarrier = SYNCS (B200_PROFILING.md 'What proves a Blackwell-native kernel')")
    for (mangled, cnt), nice in zip(per.items(), names):
        if not cnt:
            continue
        body = ", ".join(f"{k} {v}" for k, v in sorted(cnt.items()))
        print(f"{nice[:110]}\n    {body}")
