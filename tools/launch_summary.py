"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and launch count per kernel."""
import collections
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1], errors="replace")))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr = rows[hi]
    kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    acc, cnt = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) <= mv:
            continue
        name = r[kn].split("(")[0].replace("void ", "")
        try:
            v = float(r[mv].replace(",", ""))
        except ValueError:
            continue
        acc[name] += v
        cnt[name] += 1
    tot = sum(acc.values())
    print("%s: %d launches, %.3f ms" % (sys.argv[1], sum(cnt.values()), tot / 1e6))
    for k, v in acc.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 25):
        print("  %-34s %9.3f ms %5.1f%%  n=%d" % (k, v / 1e6, 100 * v / tot, cnt[k]))


if __name__ == "__main__":
    main()
