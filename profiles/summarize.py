"""Turns gpurun_out/launches_rNN.csv and the ncu reports into the tracked summaries under profiles/.
Usage: python profiles/summarize.py r1"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
out_dir = os.path.join(ROOT, "profiles")
go = os.path.join(ROOT, "gpurun_out")


def ms(v, u):
    v = float(v.replace(",", ""))
    return {"ns": v / 1e6, "us": v / 1e3, "usecond": v / 1e3, "msecond": v, "ms": v, "s": v * 1e3, "second": v * 1e3}.get(u, v)


def launches():
    rows = list(csv.reader(open(os.path.join(go, f"launches_{tag}.csv"))))
    h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H, data = rows[h], rows[h + 1:]
    ki, vi, ui, gi = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit"), H.index("Grid Size")
    agg = collections.OrderedDict()
    hist = []
    for r in data:
        if len(r) <= vi:
            continue
        name = r[ki].split("(")[0].replace("void ", "")
        t = ms(r[vi], r[ui])
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += t
        if "hist_scan" in name:
            hist.append((r[gi], t))
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(out_dir, f"{tag}_launches.md"), "w") as f:
        f.write(f"# ncu launch list, one fit of BASELINE configs[1] ({tag})\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 python profiles/profile_fit.py 1` "
                "(cold-cache, serialised: compare SHARES, not absolutes).\n\n")
        f.write("| kernel | launches | total ms | share of GPU time |\n|---|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {v[0]} | {v[1]:.3f} | {100 * v[1] / tot:.1f} % |\n")
        f.write(f"| **total** | {sum(v[0] for v in agg.values())} | {tot:.3f} | |\n\n")
        f.write("## hist_scan_kernel per launch (grid, ms)\n\n")
        f.write(", ".join(f"{g} {t:.3f}" for g, t in hist) + "\n")
    return agg, tot


def rep(name):
    p = os.path.join(go, f"{name}.ncu-rep")
    if not os.path.exists(p):
        return None
    txt = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    H, U, V = rows[0], rows[1], rows[2]
    return {h: (V[i], U[i]) for i, h in enumerate(H)}


WANT = ["gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_op_shared_atom.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum", "smsp__inst_executed.sum", "smsp__cycles_active.avg"]


def kernel_summary():
    with open(os.path.join(out_dir, f"{tag}_hist_kernel.md"), "w") as f:
        f.write(f"# hist_scan_kernel — ncu --set full ({tag})\n\n"
                "`ncu --set full --clock-control none --import-source on -k regex:hist_scan -s <n> -c 1 python profiles/profile_fit.py 1`\n"
                "Two launches of one fit of BASELINE configs[1]: launch 5 (tree level 4: 1 600 nodes, every in-bag row) and "
                "launch 21 (a deep group: 2 621 small nodes).\n\n")
        traffic = {}
        for name, label in ((f"hist_l4_{tag}", "level-4 group"), (f"hist_deep_{tag}", "deep group")):
            m = rep(name)
            if m is None:
                continue
            f.write(f"## {label} (`{name}.ncu-rep`)\n\n| metric | value | unit |\n|---|---:|---|\n")
            for w in WANT:
                if w in m:
                    f.write(f"| `{w}` | {m[w][0]} | {m[w][1]} |\n")
            f.write("\n")
            def b(k):
                v, u = m[k]
                return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
            traffic[label] = b("dram__bytes_read.sum") + b("dram__bytes_write.sum")
        return traffic


def hist_dram():
    path = os.path.join(go, f"hist_dram_{tag}.csv")
    if not os.path.exists(path):
        return None
    rows = list(csv.reader(open(path)))
    h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H, data = rows[h], rows[h + 1:]
    ii, mi, vi, ui = H.index("ID"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
    per = collections.OrderedDict()
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in data:
        if len(r) <= vi:
            continue
        d = per.setdefault(r[ii], {"bytes": 0.0, "ms": 0.0})
        if r[mi].startswith("dram__bytes"):
            d["bytes"] += float(r[vi].replace(",", "")) * scale.get(r[ui], 1)
        elif r[mi].startswith("gpu__time"):
            d["ms"] = ms(r[vi], r[ui])
    vals = list(per.values())
    tot_b, tot_ms = sum(v["bytes"] for v in vals), sum(v["ms"] for v in vals)
    out = {"kernel": "hist_scan_kernel", "launches": len(vals), "dram_bytes_per_launch": tot_b / len(vals),
           "dram_bytes_per_fit": tot_b, "kernel_ms_per_fit_under_ncu": tot_ms,
           "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:hist_scan, "
                  "every launch of one fit of BASELINE configs[1]; mean over the launches"}
    json.dump(out, open(os.path.join(out_dir, "hist_kernel_traffic.json"), "w"), indent=1)
    with open(os.path.join(out_dir, f"{tag}_hist_kernel.md"), "a") as f:
        f.write(f"## DRAM traffic of every launch in one fit\n\n{len(vals)} launches, {tot_b / 1e9:.2f} GB read+written in total "
                f"({tot_b / len(vals) / 1e6:.1f} MB per launch) against 13.36 GB of algorithmic bytes per fit "
                f"(`emrf_timings_t.hist_alg_bytes`).\n\n| launch | ms | DRAM MB |\n|---:|---:|---:|\n")
        for i, v in enumerate(vals):
            f.write(f"| {i + 1} | {v['ms']:.3f} | {v['bytes'] / 1e6:.1f} |\n")
    return out


agg, tot = launches()
traffic = kernel_summary()
print("hist dram:", hist_dram())
print("kernels:", {k: round(v[1], 3) for k, v in agg.items()}, "total", round(tot, 3))
print("traffic:", traffic)
