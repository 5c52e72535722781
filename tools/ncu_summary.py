#!/usr/bin/env python
"""Summaries of ncu exports for profiles/:

    python tools/ncu_summary.py launches <launch list .csv> "<command>" > profiles/<name>.md
    python tools/ncu_summary.py full <ncu --page raw --csv export> "<command>" > profiles/<name>.md
    python tools/ncu_summary.py traffic <ncu --page raw --csv export> <kernel key> "<source note>"   (updates profiles/r2_traffic.json)

launches: the `--metrics gpu__time_duration.sum --clock-control none --csv` pass (per-launch times are cold-cache and
serialised: shares are what matters).  full: selected metrics of every captured launch of an `ncu --set full` report.
"""
import collections
import csv
import re
import sys


def launches(path, cmd):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    tot, n = 0.0, 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1000 if row["Metric Unit"] == "ns" else v * 1000 if row["Metric Unit"] == "ms" else v
        name = re.sub(r"\(.*", "", row["Kernel Name"])[:70] + " grid=" + row["Grid Size"] + " block=" + row["Block Size"]
        a = agg.setdefault(name, [0, 0.0, []])
        a[0] += 1; a[1] += v; a[2].append(v)
        tot += v; n += 1
    print(f"# ncu launch list: `{cmd}`")
    print("# command: ncu --metrics gpu__time_duration.sum --clock-control none --csv ...")
    print("# Per-launch times are cold-cache (ncu flushes L2 between kernels) and serialised; shares are what matters.")
    print(f"# launches captured: {n}  sum of durations: {tot / 1000:.2f} ms")
    print("kernel | launches | total_ms | share | mean_us | median_us")
    for k, (c, t, l) in sorted(agg.items(), key=lambda x: -x[1][1]):
        l.sort()
        print(f"{k} | {c} | {t / 1000:.3f} | {t / tot:.3f} | {t / c:.2f} | {l[len(l) // 2]:.2f}")


FULL = [
    ("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("launch__registers_per_thread", "regs/thread"), ("launch__occupancy_limit_shared_mem", "occ limit smem (CTAs)"),
    ("launch__occupancy_limit_registers", "occ limit regs (CTAs)"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("smsp__inst_executed.sum", "warp instructions"),
]


def full(path, cmd):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    print(f"# ncu --set full capture: `{cmd}`")
    print("# command: ncu --set full --clock-control none --import-source on -k regex:... ; exported with ncu -i ... --page raw --csv")
    print("# (L2 is flushed before every profiled launch: small kernels that read another kernel's output see DRAM latency here)")
    cols = [(hdr.index(k), lab) for k, lab in FULL if k in hdr]
    ik, ig, ib = hdr.index("Kernel Name"), hdr.index("Grid Size"), hdr.index("Block Size")
    print("kernel | grid | block | " + " | ".join(lab for _, lab in cols))
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[ik])[:48]
        vals = []
        for i, _ in cols:
            try:
                vals.append(f"{float(r[i].replace(',', '')):.4g} {units[i]}".strip())
            except ValueError:
                vals.append(r[i])
        print(f"{name} | {r[ig]} | {r[ib]} | " + " | ".join(vals))


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] in ("launches", "full"):
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")


def traffic(path, key, note):
    """dram__bytes_read.sum + dram__bytes_write.sum of the first captured launch -> profiles/r2_traffic.json[key] (bench.py reads it)"""
    import json
    import os
    rows = list(csv.reader(open(path)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    get = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    tot = 0.0
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        v, u = get[name]
        tot += float(v.replace(",", "")) * mult[u]
    dur_v, dur_u = get["gpu__time_duration.sum"]
    out = "profiles/r2_traffic.json"
    d = json.load(open(out)) if os.path.exists(out) else {}
    d[key] = {"dram_bytes_per_launch": tot, "gpu_time_duration": f"{dur_v} {dur_u}", "source": note}
    json.dump(d, open(out, "w"), indent=1)
    print(key, d[key])


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "traffic":
    traffic(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
