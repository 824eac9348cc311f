"""Join an ncu launch list (gpu__time_duration per kernel) with the GEMM shape log written by BDL_LOG_GEMM to get
per-layer TFLOP/s and bytes/s of the tcgen05 kernel inside a real training step.
    BDL_LOG_GEMM=gpurun_out/shapes.jsonl ncu --metrics gpu__time_duration.sum --csv --log-file gpurun_out/l.csv python bench.py --steps 1 --warmup 1 --no-graph --no-e2e
    python benchmarks/gemm_roofline.py gpurun_out/l.csv gpurun_out/shapes.jsonl"""
import csv
import json
import sys


def main(csv_path, shapes_path, top=25):
    rows = list(csv.DictReader(l for l in open(csv_path) if not l.startswith("==")))
    times = [float(r["Metric Value"].replace(",", "")) / 1e3 for r in rows if "gemm_bf16_tcgen05" in r["Kernel Name"]]
    shapes = [json.loads(l) for l in open(shapes_path)]
    n = min(len(times), len(shapes))
    times, shapes = times[-n:], shapes[-n:]
    per_step = n // 2 if n % 2 == 0 else n          # warm-up step + timed step were both logged
    times, shapes = times[-per_step:], shapes[-per_step:]
    out = []
    for t, s in zip(times, shapes):
        if s["kind"] == "conv":
            M = s["N"] * s["Ho"] * s["Wo"]
            K = s["kh"] * s["kw"] * s["C"]
            if s["mode"] == 1:                      # dgrad: geometry given for the transposed conv
                M = s["N"] * s["Ho"] * s["Wo"]
            flops = 2.0 * M * s["O"] * K
            desc = f"conv{s['mode']} {s['C']}->{s['O']} k{s['kh']} {s['Ho']}x{s['Wo']}"
            byt = 2.0 * (M * s["C"] + M * s["O"] + s["O"] * K)
        else:
            flops = 2.0 * s["M"] * s["N"] * s["K"]
            desc = f"gemm M{s['M']} N{s['N']} K{s['K']} amn{s['a_mn']} bmn{s['b_mn']}"
            byt = 2.0 * (s["M"] * s["K"] + s["N"] * s["K"]) + (4.0 if s["acc"] else 2.0) * s["M"] * s["N"]
        out.append((t, desc, flops / t / 1e6, byt / t / 1e3))
    tot = sum(o[0] for o in out)
    print(f"{len(out)} GEMM launches, {tot / 1e3:.2f} ms")
    for t, d, tf, gb in sorted(out, reverse=True)[:top]:
        print(f"{t:8.1f} us  {tf:7.1f} TFLOP/s  {gb:7.0f} GB/s(min traffic)  {d}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
