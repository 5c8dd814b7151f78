"""Extract the judged metrics from an .ncu-rep into a small CSV and a roofline table (run where ncu is installed).

    python benchmarks/tools/ncu_extract.py prof.ncu-rep out.csv [out.md]

The markdown table aggregates by kernel name: launches, total and mean device time, achieved HBM bandwidth as a
fraction of the MEASURED copy bandwidth (MEASURED_PEAKS.json: hbm_gbs), tensor-pipe activity, L2 hit rate, occupancy.
"""
import collections
import csv
import json
import os
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
md = sys.argv[3] if len(sys.argv) > 3 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keep = ["ID", "Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.sum", "sm__inst_executed_pipe_tensor.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__inst_executed.sum",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
idx = [hdr.index(k) for k in keep if k in hdr]
with open(out, "w", newline="") as fh:
    w = csv.writer(fh)
    w.writerow([hdr[i] for i in idx])
    w.writerow([units[i] for i in idx])
    for r in rows[2:]:
        w.writerow([r[i] for i in idx])
print(len(rows) - 2, "kernels ->", out)

if md:
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    tag = "of measured" if peaks else "of fallback"

    def col(name):
        return hdr.index(name) if name in hdr else None

    def num(r, i):
        try:
            return float(r[i].replace(",", "")) if i is not None else 0.0
        except ValueError:
            return 0.0

    def scale(i, target):
        """ncu prints units per column (ns/us/ms, byte/Kbyte/...): normalise to `target` (ns or byte)."""
        u = units[i].lower() if i is not None else ""
        table = {"ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9, "second": 1e9,
                 "nsecond": 1.0, "byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}
        return table.get(u, 1.0)

    ci = {k: col(k) for k in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                              "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
                              "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size",
                              "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic")}
    agg = collections.OrderedDict()
    for r in rows[2:]:
        name = r[ci["Kernel Name"]].split("(")[0][:70]
        a = agg.setdefault(name, dict(n=0, ns=0.0, rd=0.0, wr=0.0, tensor=0.0, l2=0.0, occ=0.0, grid=0, regs=0, smem=0))
        t = num(r, ci["gpu__time_duration.sum"]) * scale(ci["gpu__time_duration.sum"], "ns")
        a["n"] += 1
        a["ns"] += t
        a["rd"] += num(r, ci["dram__bytes_read.sum"]) * scale(ci["dram__bytes_read.sum"], "byte")
        a["wr"] += num(r, ci["dram__bytes_write.sum"]) * scale(ci["dram__bytes_write.sum"], "byte")
        a["tensor"] += num(r, ci["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]) * t
        a["l2"] += num(r, ci["lts__t_sector_hit_rate.pct"]) * t
        a["occ"] += num(r, ci["sm__warps_active.avg.pct_of_peak_sustained_active"]) * t
        a["grid"] = max(a["grid"], int(num(r, ci["launch__grid_size"])))
        a["regs"] = int(num(r, ci["launch__registers_per_thread"]))
        a["smem"] = max(a["smem"], int(num(r, ci["launch__shared_mem_per_block_dynamic"]) *
                                       scale(ci["launch__shared_mem_per_block_dynamic"], "byte")))
    total = sum(a["ns"] for a in agg.values())
    with open(md, "w") as fh:
        fh.write(f"| kernel | launches | total us | mean us | share | HBM GB/s | HBM {tag} ({hbm:.0f} GB/s) | tensor pipe % | "
                 "L2 hit % | warps active % | max grid | regs | dyn smem |\n|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
            gbs = (a["rd"] + a["wr"]) / max(a["ns"], 1e-9)
            t = max(a["ns"], 1e-9)
            fh.write(f"| `{name}` | {a['n']} | {a['ns'] / 1e3:.1f} | {a['ns'] / 1e3 / a['n']:.2f} | {100 * a['ns'] / total:.1f}% | "
                     f"{gbs:.0f} | {gbs / hbm:.3f} | {a['tensor'] / t:.1f} | {a['l2'] / t:.1f} | {a['occ'] / t:.1f} | "
                     f"{a['grid']} | {a['regs']} | {a['smem']} |\n")
        fh.write(f"\nTotal device time of the captured launches: {total / 1e3:.1f} us (serialised, cold-ish caches under ncu: "
                 "compare shares and fractions, not the absolute sum).\n")
    print("roofline table ->", md)
