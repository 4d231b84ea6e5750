"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total us, share.
Usage: python tools/launch_summary.py launches.csv [out.md]"""
import csv
import re
import sys
from collections import defaultdict


def main():
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 8]
    hdr = rows[0]
    ik, im, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot = defaultdict(float)
    cnt = defaultdict(int)
    for r in rows[1:]:
        if r[im] != "gpu__time_duration.sum":
            continue
        v = float(r[iv].replace(",", ""))
        if r[iu] == "ns":
            v /= 1e3
        elif r[iu] == "ms":
            v *= 1e3
        name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").strip()
        name = re.sub(r"<.*", "", name) if len(name) > 60 else name
        tot[name] += v
        cnt[name] += 1
    total = sum(tot.values())
    lines = ["| kernel | launches | total us | avg us | share |", "|---|---:|---:|---:|---:|"]
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        lines.append(f"| {k} | {cnt[k]} | {v:.1f} | {v / cnt[k]:.2f} | {100 * v / total:.1f}% |")
    lines.append(f"| **all** | {sum(cnt.values())} | {total:.1f} | | 100% |")
    text = "\n".join(lines)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "a").write(text + "\n")


if __name__ == "__main__":
    main()
