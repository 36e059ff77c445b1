"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: the last full
training/inference step, per kernel, with shares (absolute times are cold-cache, serialised)."""
import csv
import re
import sys


def main(path, detail=False, marker="sgd_momentum"):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H = rows[hdr]
    ki, vi, gi = H.index("Kernel Name"), H.index("Metric Value"), H.index("Grid Size")
    R = [r for r in rows[hdr + 2:] if len(r) > vi]
    names = [r[ki] for r in R]
    marks = [i for i, n in enumerate(names) if marker in n]
    a, b = (marks[-2] + 1, marks[-1] + 1) if len(marks) >= 2 else (0, len(R))
    agg, tot = {}, 0.0
    for r in R[a:b]:
        n, v = r[ki], float(r[vi].replace(",", "")) / 1e3
        m = re.search(r"(conv_tc_kernel|wgrad_tc_kernel)<([^>]*)>", n)
        short = (m.group(1).replace("_kernel", "") + "<" + m.group(2) + ">") if m else n.split("(")[0].replace("canb::", "")[:40]
        tot += v
        agg.setdefault(short, [0, 0.0])
        agg[short][0] += 1
        agg[short][1] += v
        if detail and v > 30:
            print(f"{v:8.1f} us  {r[gi]:>16}  {short}")
    print(f"step: {tot:.1f} us over {b - a} launches")
    for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
        print(f"{v:9.1f} us {n:4d} {100 * v / tot:5.1f}%  {k}")


if __name__ == "__main__":
    main(sys.argv[1], detail="--detail" in sys.argv, marker=(sys.argv[sys.argv.index("--marker") + 1] if "--marker" in sys.argv else "sgd_momentum"))
