"""Per-kernel SASS evidence of the Blackwell-native paths in libssb200.so (no GPU needed):

    python tools/sass_summary.py [> profiles/sass_summary.txt]

Disassembles the in-tree library with cuobjdump and counts, per kernel, the mnemonics that prove each mechanism
(/opt/skills/guides/B200_PROFILING.md): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk (TMA, 1-D),
UTMALDG/UTMASTG = cp.async.bulk.tensor, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, LDGSTS = cp.async, HMMA = legacy
mma.sync (must be 0), plus FFMA / LDS / STS for scale and the register count from the ELF."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "silent_speech_b200", "libssb200.so")
COLS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "HMMA", "FFMA", "LDS", "STS", "SHFL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1)
            for c in COLS:
                if op == c or op.startswith(c + ".") or (c == "UTCHMMA" and op.startswith("UTC") and "MMA" in op):
                    counts[cur][c] += 1
    regs = {}
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    fn = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            fn = m.group(1)
        m = re.search(r"REG:(\d+)", line)
        if m and fn:
            regs[fn] = int(m.group(1)); fn = None
    names = demangle(list(counts))
    print("libssb200.so -- SASS mnemonics per kernel (sm_100a); tools/sass_summary.py")
    print(f"{'kernel':78s} {'REG':>4s} " + " ".join(f"{c:>7s}" for c in COLS))
    tot = collections.Counter()
    for k, c in counts.items():
        short = re.sub(r"\(.*", "", names.get(k, k)).replace("void ", "").replace("ssb::", "")
        print(f"{short[:78]:78s} {regs.get(k, 0):4d} " + " ".join(f"{c[x]:7d}" for x in COLS))
        tot.update(c)
    print(f"{'TOTAL':78s} {'':4s} " + " ".join(f"{tot[x]:7d}" for x in COLS))
    if tot["HMMA"]:
        sys.exit("legacy HMMA found")


if __name__ == "__main__":
    main()
