"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: time per kernel family and its share."""
import csv
import re
import sys
from collections import defaultdict


def family(name: str) -> str:
    m = re.search(r"(Epi\w+|AConv\w*)", name)
    base = re.sub(r"<.*", "", name.split("(")[0]).replace("zb::", "")
    return f"{base}[{m.group(1)}]" if m and "gemm" in base else base


def main(path, frames):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("==")) if r]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(r[ui], 1e-3)
        f = family(r[ki])
        tot[f] += v * scale
        cnt[f] += 1
    total = sum(tot.values())
    print(f"{path}: {sum(cnt.values())} launches over {frames} frames, {total / frames:.1f} us of kernel time per frame")
    print(f"{'kernel':58s} {'n/frame':>8s} {'us/launch':>10s} {'us/frame':>10s} {'share':>7s}")
    for f in sorted(tot, key=tot.get, reverse=True):
        print(f"{f:58s} {cnt[f] / frames:8.1f} {tot[f] / cnt[f]:10.2f} {tot[f] / frames:10.1f} {100 * tot[f] / total:6.1f}%")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 1)
