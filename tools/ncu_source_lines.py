"""Condense an `ncu -i rep --page source --csv --print-source cuda,sass` export into the hottest source lines per kernel:
share of executed warp instructions and of stall samples.  usage: python tools/ncu_source_lines.py <export.csv> <out.md> [title] [top]"""
import csv
import sys


def main():
    src, out = sys.argv[1], sys.argv[2]
    title = sys.argv[3] if len(sys.argv) > 3 else src
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(src, errors="replace")))
    cur = func = None
    agg = {}
    for r in rows:
        if len(r) >= 2 and r[0] == "Function Name":
            func = r[1]; continue
        if len(r) >= 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]; continue
        if len(r) > 8 and r[0].isdigit() and r[2] == "-":
            try:
                n, s = int(r[7]), int(r[6])
            except ValueError:
                continue
            a = agg.setdefault(func, {}).setdefault((cur, int(r[0])), [0, 0, r[1]])
            a[0] += n; a[1] += s
    lines = ["# " + title, "", "From `ncu --set full --import-source on` (source page, CUDA view): per source line the share of executed warp instructions and of "
             "warp-stall samples of the kernel.", ""]
    for f, d in agg.items():
        tot = sum(a[0] for a in d.values()); ts = sum(a[1] for a in d.values())
        if tot == 0:
            continue
        lines += ["## `%s`" % f, "", "%d warp instructions, %d stall samples" % (tot, ts), "", "| inst % | cumulative | samples % | line | source |", "|---|---|---|---|---|"]
        acc = 0
        for (fl, l), (n, s, text) in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
            acc += n
            lines.append("| %.1f | %.1f | %.1f | %s:%d | `%s` |" % (100 * n / tot, 100 * acc / tot, 100 * s / max(ts, 1), fl, l, text.strip().replace("|", "\\|")[:110]))
        lines.append("")
    open(out, "w").write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
