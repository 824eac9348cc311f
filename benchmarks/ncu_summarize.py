"""Condense .ncu-rep files into one table (kernel, time, DRAM / L2 / SM utilisation, occupancy, registers, top stalls):
    python benchmarks/ncu_summarize.py gpurun_out/ncu_*.ncu-rep > profiles/ncu_kernels_r2.md"""
import csv, io, subprocess, sys

KEYS = {
    "time_us": "gpu__time_duration.sum",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram_GBs": "dram__bytes.sum.per_second",
    "l2_pct": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm_pct": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "issue_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "occ_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "tensor_pct": "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "regs": "launch__registers_per_thread",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
}


def rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr, units = r[0], r[1]
    for row in r[2:]:
        d = dict(zip(hdr, row))
        u = dict(zip(hdr, units))
        rec = {"kernel": d.get("Kernel Name", "?")}
        for k, m in KEYS.items():
            col = next((h for h in hdr if h == m or h.endswith("." + m) or h.endswith(m)), None)
            v = d.get(col, "") if col else ""
            try:
                v = float(v.replace(",", ""))
                if k == "time_us":
                    unit = u.get(col, "")
                    v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(unit, 1)
                if k == "dram_GBs":
                    unit = u.get(col, "")
                    v *= {"byte/s": 1e-9, "Kbyte/s": 1e-6, "Mbyte/s": 1e-3, "Gbyte/s": 1, "Tbyte/s": 1e3}.get(unit, 1e-9)
            except ValueError:
                v = None
            rec[k] = v
        st = []
        for h in hdr:
            if "smsp__average_warps_issue_stalled_" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
                try:
                    st.append((float(d[h]), h.split("stalled_")[1].replace("_per_issue_active.ratio", "")))
                except ValueError:
                    pass
        st.sort(reverse=True)
        rec["stalls"] = ", ".join(f"{n} {v:.1f}" for v, n in st[:3] if n != "selected")
        yield rec


def fmt(v, nd=1):
    return "-" if v is None else (f"{v:.{nd}f}" if isinstance(v, float) else str(v))


print("| report | kernel | time us | DRAM % | DRAM GB/s | L2 % | SM % | issue % | warps active % | tensor % | regs | grid x block | top stalls (warps per issue) |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for rep in sys.argv[1:]:
    seen = {}
    for r in rows(rep):
        name = r["kernel"].replace("void ", "").replace("bdl::", "")
        name = name[:70]
        key = (name, r["grid"])
        if key in seen:
            continue
        seen[key] = 1
        print(f"| {rep.split('/')[-1].replace('.ncu-rep', '')} | `{name}` | {fmt(r['time_us'])} | {fmt(r['dram_pct'])} | {fmt(r['dram_GBs'], 0)} | "
              f"{fmt(r['l2_pct'])} | {fmt(r['sm_pct'])} | {fmt(r['issue_pct'])} | {fmt(r['occ_pct'])} | {fmt(r['tensor_pct'])} | "
              f"{fmt(r['regs'], 0)} | {fmt(r['grid'], 0)} x {fmt(r['block'], 0)} | {r['stalls']} |")
