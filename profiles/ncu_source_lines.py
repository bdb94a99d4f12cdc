"""Top source lines by executed warp instructions / stall samples from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv`."""
import csv
import sys


def main(path, top=40):
    cur, hdr, out = "?", None, []
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] in ("File Path", "File Name"):
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit() and len(r) > 8 and r[2] == "-":
            i, s = hdr.index("Instructions Executed"), hdr.index("# Samples")
            try:
                out.append((int(r[i]), int(r[s]), float(r[hdr.index("Avg. Threads Executed")]), cur, int(r[0]), r[1].strip()[:100]))
            except ValueError:
                pass
    ti, ts = sum(o[0] for o in out), sum(o[1] for o in out)
    print("total warp-instructions %d, samples %d" % (ti, ts))
    print("%7s %7s %5s  %s" % ("inst%", "smpl%", "thr", "line"))
    for o in sorted(out, key=lambda o: -o[1])[:top]:
        print("%6.2f%% %6.2f%% %5.1f  %s:%d  %s" % (100.0 * o[0] / ti, 100.0 * o[1] / max(ts, 1), o[2], o[3], o[4], o[5]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
