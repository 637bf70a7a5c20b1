"""Per-kernel histogram of the SASS mnemonics that prove which hardware paths the shipped libraries use:
    python profiles/make_sass_summary.py > profiles/sass_summary.txt
DMMA = fp64 tensor pipe (mma.sync m8n8k4.f64), LDGSTS = cp.async, UBLKCP = cp.async.bulk (1-D TMA), UTMALDG =
cp.async.bulk.tensor (tiled TMA through a tensor map), SYNCS = mbarrier operations, UCGABAR = cluster barrier,
UTCIMMA / LDTM = tcgen05 (tcgen05.mma kind::i8 / tcgen05.ld: the Ozaki-split int8 GEMM of csrc/ozaki_gemm.cuh; tcgen05 has no
f64 kind, so the fp64 kernels carry none)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["DMMA", "DFMA", "LDGSTS", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "UCGABAR", "BAR.SYNC", "UTCIMMA", "LDTM", "LDS", "STS"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    for lib in ("draco_b200/libdgibbs.so", "draco_b200/libdgnrml.so"):
        sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, lib)], capture_output=True, text=True).stdout
        counts, cur = collections.OrderedDict(), None
        for line in sass.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                cur = m.group(1)
                counts[cur] = collections.Counter()
                continue
            if cur is None:
                continue
            m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if m:
                op = m.group(1)
                counts[cur]["total"] += 1
                for k in KEYS:
                    if op.startswith(k) and not (k == "LDS" and op.startswith("LDSM")):
                        counts[cur][k] += 1
        names = demangle(list(counts))
        print(f"== {lib}  (cuobjdump -sass, sm_100a)")
        print(f"{'kernel':72s} {'instr':>7s} " + " ".join(f"{k:>8s}" for k in KEYS))
        tot = collections.Counter()
        for fn, c in counts.items():
            nm = re.sub(r"\(.*", "", names.get(fn, fn))[:72]
            print(f"{nm:72s} {c['total']:7d} " + " ".join(f"{c[k]:8d}" for k in KEYS))
            tot.update(c)
        print(f"{'TOTAL':72s} {tot['total']:7d} " + " ".join(f"{tot[k]:8d}" for k in KEYS))
        print()


if __name__ == "__main__":
    sys.exit(main())
