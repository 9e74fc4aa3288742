"""cuobjdump -sass of libb200approx.so -> per-kernel counts of the Blackwell-specific opcodes
(B200_PROFILING.md: UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTCBAR = tcgen05.commit,
UBLKCP = cp.async.bulk on the TMA engine, SYNCS = mbarrier).
    python profiles/sass_summary.py > profiles/r02/sass_summary.txt
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "util_b200", "csrc", "libb200approx.so")
WATCH = ["UTCHMMA", "UTCQMMA", "UTCIMMA", "LDTM", "STTM", "UTCBAR", "UTCCP", "UBLKCP", "UTMALDG", "SYNCS", "FMNMX3",
         "DFMA", "DADD", "DMUL", "HMMA", "MEMBAR", "NANOSLEEP", "LDL", "STL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    per = collections.OrderedDict()
    name = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", ""))
            per[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", line)
        if m and name:
            per[name][m.group(1).split(".")[0]] += 1
            per[name]["_total"] += 1
    print(f"{os.path.relpath(LIB, ROOT)}: {len(per)} kernels, sm_100a SASS")
    print(f"{'kernel':70s} {'instr':>7s}  " + " ".join(f"{w:>8s}" for w in WATCH))
    for k, c in per.items():
        if not any(c[w] for w in WATCH[:11]) and "knn" not in k and "lshep" not in k and "shepmod" not in k:
            continue
        print(f"{k[-70:]:70s} {c['_total']:7d}  " + " ".join(f"{c[w]:8d}" for w in WATCH))
    tot = collections.Counter()
    for c in per.values():
        tot.update(c)
    print(f"{'ALL KERNELS':70s} {tot['_total']:7d}  " + " ".join(f"{tot[w]:8d}" for w in WATCH))


if __name__ == "__main__":
    main()
