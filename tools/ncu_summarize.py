"""Turn `ncu -i X.ncu-rep --page raw --csv` output into the markdown summary + traffic JSON kept under profiles/.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_summarize.py raw.csv "title line" profiles/rX_ncu_full_summary.md profiles/rX_traffic.json BYTES_PER_LAUNCH
"""
import csv, json, re, sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__cycles_active.avg", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__waves_per_multiprocessor", "sm__cycles_elapsed.avg.per_second", "dram__cycles_elapsed.avg.per_second",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    raw, title, md_path, json_path, per_launch = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4], int(sys.argv[5])
    rows = list(csv.reader(open(raw)))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    head, units, data = rows[start], rows[start + 1], rows[start + 2:]
    col = {n: i for i, n in enumerate(head)}
    out, traffic = ["# " + title, ""], {"source": "ncu --set full capture %s" % md_path}
    for r in data:
        if len(r) < len(head):
            continue
        name = r[col["Kernel Name"]]
        out += ["## " + name, "", "| metric | value | unit |", "|---|---|---|"]
        for m in KEEP:
            if m in col:
                out.append("| %s | %s | %s |" % (m, r[col[m]], units[col[m]]))
        out.append("")
        m = re.search(r"(\w+)(?:<[^<>()]*>)?\(", name)  # "void <unnamed>::hbm_copy_kernel<1>(char *, ...)" -> hbm_copy_kernel
        key = m.group(1) if m else name
        try:
            rd = float(r[col["dram__bytes_read.sum"]].replace(",", "")) * UNIT_SCALE[units[col["dram__bytes_read.sum"]]]
            wr = float(r[col["dram__bytes_write.sum"]].replace(",", "")) * UNIT_SCALE[units[col["dram__bytes_write.sum"]]]
            traffic[key] = {"bytes_per_launch": per_launch, "dram_bytes_read": int(rd), "dram_bytes_write": int(wr)}  # last launch of each kernel wins
        except (KeyError, ValueError):
            pass
    open(md_path, "w").write("\n".join(out) + "\n")
    open(json_path, "w").write(json.dumps(traffic, indent=1) + "\n")


def launches():
    """python tools/ncu_summarize.py --launches launches.csv out.txt : per-kernel count / total / share of a
    `--metrics gpu__time_duration.sum --csv` launch list."""
    rows = list(csv.reader(open(sys.argv[2], errors="replace")))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    col = {n: i for i, n in enumerate(rows[start])}
    agg, order = {}, []
    for r in rows[start + 1:]:
        if len(r) <= col["Metric Value"] or r[col["Metric Name"]] != "gpu__time_duration.sum":
            continue
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}[r[col["Metric Unit"]]]
        us = float(r[col["Metric Value"]].replace(",", "")) * scale
        k = r[col["Kernel Name"]]
        if k not in agg:
            agg[k] = [0, 0.0]
            order.append(k)
        agg[k][0] += 1
        agg[k][1] += us
    total = sum(v[1] for v in agg.values())
    with open(sys.argv[3], "w") as f:
        for k in order:
            n, us = agg[k]
            f.write("%-78s n=%3d total=%10.1f us avg=%9.1f us share=%5.1f%%\n" % (k[:78], n, us, us / n, 100 * us / total))
        f.write("TOTAL GPU time in kernels: %.1f us over %d launches\n" % (total, sum(v[0] for v in agg.values())))


if __name__ == "__main__":
    launches() if sys.argv[1] == "--launches" else main()
