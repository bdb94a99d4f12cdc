"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count / mean / total / share."""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h = rows[hdr]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.defaultdict(list)
    for r in rows[hdr + 1:]:
        if len(r) > vi:
            agg[r[ki].split("(")[0][:70]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    print("%-70s %5s %12s %12s %6s" % ("kernel", "n", "mean_us", "total_us", "share"))
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("%-70s %5d %12.2f %12.2f %5.1f%%" % (k, len(v), sum(v) / len(v) / 1e3, sum(v) / 1e3, 100 * sum(v) / tot))


if __name__ == "__main__":
    main(sys.argv[1])
