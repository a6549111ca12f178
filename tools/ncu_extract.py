"""Turn an ncu report (--set full) into a markdown summary: one table row per kernel launch with the metrics the
roofline discussion needs.  Runs `ncu -i <rep> --page raw --csv` (so it works on the GPU box and here).
usage: python tools/ncu_extract.py <rep.ncu-rep> <out.md> [title]"""
import csv
import io
import subprocess
import sys

COLS = [
    ("gpu__time_duration.sum", "time", 1e-3, "us"),
    ("launch__grid_size", "grid", 1, ""),
    ("launch__block_size", "block", 1, ""),
    ("launch__registers_per_thread", "regs", 1, ""),
    ("dram__bytes_read.sum", "dram rd", None, "MB"),
    ("dram__bytes_write.sum", "dram wr", None, "MB"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %", 1, ""),
    ("lts__t_bytes.sum", "L2 bytes", None, "MB"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %", 1, ""),
    ("smsp__inst_executed.sum", "warp inst", 1e-6, "M"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %", 1, ""),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes/inst", 1, ""),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %", 1, ""),
]
UNIT_SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6}


def main():
    rep, out = sys.argv[1], sys.argv[2]
    title = sys.argv[3] if len(sys.argv) > 3 else rep
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    open(out.replace(".md", "") + ".raw.csv", "w").write(txt)      # kept: the summary below is a view of this file
    rows = list(csv.reader(io.StringIO(txt)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[hi], rows[hi + 1]
    kn = hdr.index("Kernel Name")
    idx = {name: hdr.index(name) for name, *_ in COLS if name in hdr}
    lines = ["# " + title, "", "Source: `ncu --set full --clock-control none --import-source on`, one row per captured launch (cold cache, "
             "serialised by the profiler).  dram rd + dram wr = the `traffic` of bench.py's roofline.", "",
             "| kernel | " + " | ".join(lbl + (" (" + u + ")" if u else "") for _, lbl, _, u in COLS) + " | DRAM GB/s |", "|---|" + "---|" * (len(COLS) + 1)]
    agg = {}
    for r in rows[hi + 2:]:
        if len(r) <= kn:
            continue
        name = r[kn].split("(")[0].replace("void ", "")
        vals, num = [], {}
        for mname, lbl, scale, u in COLS:
            if mname not in idx:
                vals.append("-"); continue
            raw = r[idx[mname]].replace(",", "")
            try:
                v = float(raw)
            except ValueError:
                vals.append(raw); continue
            if scale is None:
                v *= UNIT_SCALE.get(units[idx[mname]], 1e-6)
            elif mname == "gpu__time_duration.sum":
                v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(units[idx[mname]], 1e-3)
            else:
                v *= scale
            num[lbl] = v
            vals.append("%.3f" % v if abs(v) < 100 else "%.1f" % v)
        gbs = (num.get("dram rd", 0) + num.get("dram wr", 0)) * 1e6 / (num["time"] * 1e-6) / 1e9 if num.get("time") else 0
        lines.append("| " + name + " | " + " | ".join(vals) + " | %.0f |" % gbs)
        a = agg.setdefault(name, [0, 0.0, 0.0])
        a[0] += 1; a[1] += num.get("time", 0); a[2] += num.get("dram rd", 0) + num.get("dram wr", 0)
    lines += ["", "## per kernel", "", "| kernel | launches | total time (us) | total DRAM traffic (MB) | GB/s |", "|---|---|---|---|---|"]
    for k, (n, t, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append("| %s | %d | %.1f | %.1f | %.0f |" % (k, n, t, b, b * 1e6 / (t * 1e-6) / 1e9 if t else 0))
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[-(len(agg) + 3):]))


if __name__ == "__main__":
    main()
