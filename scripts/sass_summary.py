"""Per-kernel SASS mnemonic counts of libbaobzi.so (static instruction counts), kept under profiles/ as the evidence
for what each kernel is made of: FP64 (DFMA / DADD / DMUL / DMMA), async copies (LDGSTS = cp.async, UBLKCP = TMA bulk
copy, SYNCS = mbarrier), shared atomics, uniform-datapath loads.  usage: python scripts/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "baobzi_b200", "lib", "libbaobzi.so")
KEEP = ["DFMA", "DADD", "DMUL", "DMMA", "DSETP", "LDS", "STS", "ATOMS", "LDG", "STG", "RED", "ATOMG", "LDGSTS", "UBLKCP",
        "UBLKPF", "SYNCS", "LDC", "LDCU", "ULDC", "SHFL", "MATCH", "BAR", "F2I", "HMMA", "UTCHMMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    dem = {}
    counts = collections.OrderedDict()
    cur = None
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m and cur:
            counts[cur][m.group(1)] += 1
            counts[cur]["_total"] += 1
    names = list(counts)
    if names:
        r = subprocess.run(["c++filt"] + names, stdout=subprocess.PIPE, text=True)
        dem = dict(zip(names, r.stdout.splitlines()))
    print("# static SASS instruction counts per kernel of baobzi_b200/lib/libbaobzi.so (sm_100a); columns with a zero are left out")
    for n in sorted(names, key=lambda k: dem.get(k, k)):
        c = counts[n]
        short = re.sub(r"\(bz::TreeView.*", "", dem.get(n, n)).replace("void bz::", "")
        cols = " ".join("%s=%d" % (k, c[k]) for k in KEEP if c[k])
        print("%-58s total=%-6d %s" % (short[:58], c["_total"], cols))
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print("# library totals: " + " ".join("%s=%d" % (k, tot[k]) for k in KEEP if tot[k]))


if __name__ == "__main__":
    main()
