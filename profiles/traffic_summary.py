"""DRAM / L2 traffic per kernel from `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,
lts__t_bytes.sum --csv`: mean per launch, and the per-call sum over the rasterize+concat stage (one launch of each of
setup / raster / gt_norm / finalize_live per rac_render_concat call).  Writes JSON when given an output path.
Usage: python profiles/traffic_summary.py traffic.csv [out.json]"""
import collections
import csv
import json
import sys

STAGE = ("setup_kernel", "raster_kernel", "gt_norm_kernel", "finalize_live_kernel")


def main(path, out=None):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h = rows[hdr]
    ki, mi, vi = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
    agg = collections.defaultdict(lambda: collections.defaultdict(list))
    for r in rows[hdr + 1:]:
        if len(r) > vi:
            agg[r[ki].split("(")[0]][r[mi]].append(float(r[vi].replace(",", "")))
    res = {}
    for k, m in agg.items():
        res[k] = {"launches": len(m["gpu__time_duration.sum"]),
                  "time_us": sum(m["gpu__time_duration.sum"]) / len(m["gpu__time_duration.sum"]) / 1e3,
                  "dram_read_bytes": sum(m["dram__bytes_read.sum"]) / len(m["dram__bytes_read.sum"]),
                  "dram_write_bytes": sum(m["dram__bytes_write.sum"]) / len(m["dram__bytes_write.sum"]),
                  "l2_bytes": sum(m["lts__t_bytes.sum"]) / len(m["lts__t_bytes.sum"])}
    stage = sum(res[k]["dram_read_bytes"] + res[k]["dram_write_bytes"] for k in STAGE if k in res)
    res["stage_dram_bytes_per_call"] = stage
    for k, v in res.items():
        print(k, json.dumps(v))
    if out:
        json.dump(res, open(out, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
