"""Aggregate an ncu launch list (benchmarks/kernel_breakdown.sh) by kernel for the last of the 4 recorded steps.
usage: python benchmarks/kernel_breakdown.py gpurun_out/launches_<tag>.csv [top]"""
import csv
import re
import sys
from collections import defaultdict


def main(path, top=20, steps=4):
    rows = list(csv.DictReader(l for l in open(path) if not l.startswith("==")))
    per = len(rows) // steps
    agg = defaultdict(lambda: [0, 0.0])
    for r in rows[-per:]:
        name = r["Kernel Name"]
        k = re.sub(r"\(.*", "", name)[:72]
        if "gemm_bf16_tcgen05" in name:
            m = re.search(r"<(.*?)>", name)
            k = "gemm_bf16_tcgen05<" + (m.group(1) if m else "") + ">"
        agg[k][0] += 1
        agg[k][1] += float(r["Metric Value"].replace(",", "")) / 1e3
    tot = sum(v[1] for v in agg.values())
    print(f"{per} launches per step, {tot / 1e3:.2f} ms of kernel time (serialised, cold caches under ncu)")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{v[1]:9.1f} us {v[0]:4d} x  {k}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 20)
