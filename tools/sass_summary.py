"""Per-kernel counts of the SASS mnemonics that prove which hardware paths libpq_b200.so uses
(B200_PROFILING.md: UTCHMMA = tcgen05.mma, UTMALDG / UTMAPF = TMA load / prefetch, LDTM = tcgen05.ld,
HMMA = mma.sync, LDGSTS = cp.async, SYNCS = mbarrier).  Usage:
    python tools/sass_summary.py [path/to/lib.so] > profiles/sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MNEMONICS = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMAPF", "LDTM", "STTM", "UTCBAR", "HMMA", "LDGSTS", "SYNCS",
             "UCGABAR", "LDS", "STS", "LDG", "STG", "MEMBAR", "ATOM", "RED", "ST.E.64.STRONG.SYS", "LD.E.64.STRONG.SYS")


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vllm_pq_b200", "libpq_b200.so")
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    demangle = {}
    counts = collections.OrderedDict()
    cur = None
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
            counts[cur]["_total"] += 1
            for mn in MNEMONICS:
                if op == mn or op.startswith(mn + "."):
                    counts[cur][mn] += 1
    names = list(counts)
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    for n, d in zip(names, out):
        demangle[n] = d
    print(f"# {os.path.relpath(lib, ROOT)}: {len(names)} kernels (sm_100a SASS, cuobjdump -sass)")
    totals = collections.Counter()
    groups = collections.OrderedDict()
    for n in names:
        d = demangle.get(n, n)
        short = re.sub(r"<.*", "", d.replace("void ", "").replace("pq::", ""))
        g = groups.setdefault(short, [0, collections.Counter()])
        g[0] += 1
        g[1].update(counts[n])
        totals.update(counts[n])
    cols = [m for m in MNEMONICS if totals[m]]
    print(f"{'kernel (template instantiations)':44s} {'inst':>5s} {'SASS':>8s} " + " ".join(f"{c[:10]:>10s}" for c in cols))
    for short, (k, c) in groups.items():
        print(f"{short[:44]:44s} {k:5d} {c['_total']:8d} " + " ".join(f"{c[m]:10d}" for m in cols))
    print(f"{'total':44s} {len(names):5d} {totals['_total']:8d} " + " ".join(f"{totals[m]:10d}" for m in cols))


if __name__ == "__main__":
    main()
