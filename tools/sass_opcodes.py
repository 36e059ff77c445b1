"""Per-kernel counts of the Blackwell-specific SASS opcodes in the built library (tcgen05 MMA = UTC*MMA, TMEM loads =
LDTM, TMA loads / stores = UTMALDG / UTMASTG, tcgen05.commit = UTCBAR, mbarrier = SYNCS):

    python tools/sass_opcodes.py > profiles/r02_sass_opcodes.txt        (needs cuobjdump; no GPU)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "can-distributed-pytorch_b200", "can_b200", "libcan_b200.so")
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UTCBAR[.\w]*|UTMAPF[.\w]*|UTMACMDFLUSH|SYNCS[.\w]*)")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], check=True, capture_output=True, text=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    per, total, fn, i = collections.OrderedDict(), collections.Counter(), None, -1
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            i += 1
            fn = names[i] if i < len(names) else m.group(1)
            fn = re.sub(r"\(.*", "", fn)
            per[fn] = collections.Counter()
            continue
        m = PAT.search(line)
        if m and fn:
            op = m.group(1)
            per[fn][op] += 1
            total[op] += 1
    print(f"# {os.path.relpath(SO, ROOT)}: Blackwell opcodes per kernel (cuobjdump -sass)")
    print("# total: " + ", ".join(f"{k} {v}" for k, v in sorted(total.items())))
    multicast = sum(v for k, v in total.items() if k.startswith("UTMALDG") and "MULTICAST" in k)
    print(f"# TMA loads with multicast in the library: {multicast}  (UTCBAR...MULTICAST = tcgen05.commit arriving on both CTAs of a pair)")
    for fn, c in per.items():
        if c:
            print(f"{fn}: " + ", ".join(f"{k} {v}" for k, v in sorted(c.items())))


if __name__ == "__main__":
    sys.exit(main())
