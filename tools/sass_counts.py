"""Per-kernel counts of the Blackwell-only SASS mnemonics in libcvb200.so (cuobjdump -sass), the evidence
B200_PROFILING.md asks for: UTC*MMA = tcgen05.mma, UTMALDG/UTMASTG = TMA, LDTM/STTM = tcgen05.ld/st, UTCBAR = tcgen05.commit,
SYNCS = mbarrier, HMMA = legacy mma.sync.   python tools/sass_counts.py > profiles/r02_sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cosyvoice_b200", "lib", "libcvb200.so")
PATTERNS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "HMMA", "MUFU", "FFMA2"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for p in PATTERNS:
            if re.search(r"\b" + p + r"[A-Z0-9_.]*\b", line):
                counts[cur][p] += 1
    demangled = {}
    try:
        names = list(counts)
        d = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
        demangled = dict(zip(names, d))
    except Exception:
        pass
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} : instruction counts per kernel (sm_100a)")
    print("# " + " ".join(f"{p:>8s}" for p in PATTERNS) + "  kernel")
    for k, c in counts.items():
        if not any(c[p] for p in PATTERNS[:10]) and "--all" not in sys.argv:
            continue
        name = demangled.get(k, k)
        name = name.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        name = re.sub(r"\(.*", "", name)
        print("  " + " ".join(f"{c[p]:8d}" for p in PATTERNS) + "  " + name)


if __name__ == "__main__":
    main()
