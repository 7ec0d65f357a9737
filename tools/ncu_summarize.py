"""Turn `ncu -i X.ncu-rep --page raw --csv` output into the markdown summary + traffic JSON kept under profiles/.

    ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_summarize.py raw.csv "title line" profiles/rX_ncu_full_summary.md profiles/rX_traffic.json BYTES_PER_LAUNCH
"""
import csv, json, sys

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
        key = name.split("(")[0].split("<")[0].split("::")[-1]
        try:
            rd = float(r[col["dram__bytes_read.sum"]].replace(",", "")) * UNIT_SCALE[units[col["dram__bytes_read.sum"]]]
            wr = float(r[col["dram__bytes_write.sum"]].replace(",", "")) * UNIT_SCALE[units[col["dram__bytes_write.sum"]]]
            traffic[key] = {"bytes_per_launch": per_launch, "dram_bytes_read": int(rd), "dram_bytes_write": int(wr)}  # last launch of each kernel wins
        except (KeyError, ValueError):
            pass
    open(md_path, "w").write("\n".join(out) + "\n")
    open(json_path, "w").write(json.dumps(traffic, indent=1) + "\n")


if __name__ == "__main__":
    main()
