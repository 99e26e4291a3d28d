"""Per-kernel SASS evidence for profiles/: counts of the Blackwell mnemonics that prove what a kernel is
made of (tcgen05 MMA = UTC*MMA, TMEM loads = LDTM, TMA = UTMALDG, mbarrier-multicast commits = UTCBAR,
128-bit global accesses, global atomics / reductions, system-scope accesses for the peer exchange, and
any legacy mma.sync).  Reads `cuobjdump -sass graph_learning_b200/libgl_b200.so`; works without a GPU.

    python tools/sass_summary.py > profiles/r2_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "graph_learning_b200", "libgl_b200.so")
PATTERNS = [("UTC*MMA", r"\bUTC[HQIO]MMA"), ("2CTA", r"UTC[HQIO]MMA\.2CTA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"),
            ("UTCBAR", r"\bUTCBAR"), ("SYNCS", r"\bSYNCS"), ("LDG.128", r"\bLDG\.E(\.\w+)*\.128"),
            ("STG.128", r"\bSTG\.E(\.\w+)*\.128"), ("LDGSTS", r"\bLDGSTS"), ("ATOMG/RED", r"\b(ATOMG|RED)\."),
            ("ATOMS", r"\bATOMS"), (".SYS", r"\.SYS\b"), ("mma.sync", r"\b[HID]MMA\.")]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts = collections.OrderedDict()
    fn = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = collections.Counter()
            continue
        if fn is None or "/*" not in line:
            continue
        counts[fn]["instructions"] += 1
        for name, pat in PATTERNS:
            if re.search(pat, line):
                counts[fn][name] += 1
    names = demangle(list(counts))
    cols = [p[0] for p in PATTERNS]
    print("# cuobjdump -sass graph_learning_b200/libgl_b200.so (sm_100a), mnemonic counts per kernel")
    print("# " + " ".join(f"{c:>9s}" for c in ["instr"] + cols) + "  kernel")
    for fn in sorted(counts, key=lambda f: names[f]):
        c = counts[fn]
        short = re.sub(r"\(.*", "", names[fn])[:120]
        print("  " + " ".join(f"{c[k]:9d}" for k in ["instructions"] + cols) + "  " + short)


if __name__ == "__main__":
    sys.exit(main())
