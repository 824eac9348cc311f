"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel time of the LAST training step."""
import collections
import csv
import re
import sys


def main(path, steps_total=4, top=14):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    per = len(rows) // steps_total
    last = rows[-per:]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in last:
        name = re.sub(r"\(.*", "", re.sub(r"<.*", "", row["Kernel Name"]))[:48]
        if "gemm_bf16_tcgen05" in row["Kernel Name"]:
            m = re.search(r"<\(bool\)(\d), \(bool\)(\d), \(int\)(\d)>", row["Kernel Name"])
            if m:
                name = "gemm_tcgen05<A%s,B%s,conv%s>" % m.groups()
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else (v * 1e3 if row["Metric Unit"] == "ms" else v)
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{path}: launches/step={per} total={tot/1e3:.2f} ms")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"{v[1]:9.1f} us {100*v[1]/tot:5.1f}%  n={v[0]:4d}  {k}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 4)
