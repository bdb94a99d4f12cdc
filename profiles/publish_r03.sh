#!/bin/bash
# gpurun_out/r03 (written by profiles/collect_r03.sh on the GPU box) -> the committed profiles/*_r03* files.  Run here (no GPU).
set -eu
R=gpurun_out/r03
cp $R/bench.json profiles/bench_r03.json
cp $R/bench_ref.json profiles/bench_ref_r03.json
cp $R/e2e_breakdown_driver.txt profiles/e2e_breakdown_r03_driver.txt
cp $R/e2e_breakdown_u8.txt profiles/e2e_breakdown_r03_u8.txt
cp $R/gt_norm_bench.txt profiles/gt_norm_bench_r03.txt
cp $R/launches.csv profiles/launches_r03.csv
python profiles/summarize_launches.py $R/launches.csv > profiles/launches_r03_summary.txt
{ echo; echo "librac kernels only (one serial-chain step = memset + setup + raster + gt_norm (forked) + finalize_live + pose_loss_step):"; grep "rac::" profiles/launches_r03_summary.txt; } > /tmp/rac_only.txt
cat /tmp/rac_only.txt >> profiles/launches_r03_summary.txt
ncu -i $R/stage.ncu-rep --page raw --csv > /tmp/stage_raw.csv 2>/dev/null
python profiles/ncu_raw_summary.py /tmp/stage_raw.csv > profiles/kernels_r03_raw.txt
ncu -i $R/assemble.ncu-rep --page raw --csv > /tmp/asm_raw.csv 2>/dev/null
python profiles/ncu_raw_summary.py /tmp/asm_raw.csv > profiles/assemble_r03_raw.txt
ncu -i $R/stage.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:raster_kernel > /tmp/raster_src.csv 2>/dev/null
python profiles/ncu_source_lines.py /tmp/raster_src.csv 40 > profiles/raster_r03_lines.txt
python profiles/traffic_summary.py $R/traffic.csv /tmp/traffic.json > /dev/null
python - <<'PY'
import csv, json, statistics, sys, importlib.util
sys.argv = ['x']
spec = importlib.util.spec_from_file_location("bench", "bench.py"); b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
t = json.load(open('/tmp/traffic.json'))
t["lib_digest"] = b.stage_source_digest()
t["note"] = ("ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum over 12 calls of the stage inside "
             "`python bench.py --steps 3 --warmup 3 --no-cpu-baseline` (profiles/collect_r03.sh); lib_digest = bench.stage_source_digest(): sha256 over the "
             "stage's kernel sources (raster*.cu/.cuh/.inl, rac_common.cuh) of the tree the capture ran on")
json.dump(t, open('profiles/traffic_r03.json', 'w'), indent=1)
rows = list(csv.reader(open('gpurun_out/r03/clocks.csv')))
hdr = [h.strip() for h in rows[0]]
ci = [i for i, h in enumerate(hdr) if h.startswith('clocks.current.sm')][0]
sm = [int(r[ci].split()[0]) for r in rows[1:] if len(r) > ci and r[ci].split()[0].isdigit()]
busy = [v for v in sm if v > 1000]
ri = [i for i, h in enumerate(hdr) if 'clocks_event_reasons.' in h and 'active' not in h]
act = sorted({tuple(r[i].strip() for i in ri) for r in rows[1:] if len(r) > max(ri)})
open('profiles/clocks_r03.txt', 'w').write(
    "nvidia-smi -lms 200 during `bench.py --impl reference` (GPU idle) and `bench.py` (profiles/collect_r03.sh): %d samples, %d under load with SM clock "
    "median %s MHz (min %s, max %s of 1965); %s -> distinct rows seen: %s\n" % (len(sm), len(busy), statistics.median(busy) if busy else None,
    min(busy) if busy else None, max(busy) if busy else None, [hdr[i].replace('clocks_event_reasons.', '') for i in ri], act))
d = json.load(open('profiles/bench_r03.json')); r = json.load(open('profiles/bench_ref_r03.json'))
print("value %.0f  ms/step %.4f  overlapped %.0f  it/s+cnn %.2f  stage_ms %.4f frac %.4f traffic %s" % (d["value"], d["ms_per_step"], d["value_overlapped"], d["iters_per_sec_with_cnn"], d["roofline"]["stage_ms"], d["roofline"]["frac"], d["roofline"]["traffic"]))
for k in ("e2e", "e2e_fresh_gt", "e2e_pinned_f32", "e2e_u8", "e2e_batch"):
    print(k, "%.0f" % d[k]["value"], d[k].get("ms_per_step", d[k].get("ms_per_batch")))
print("cpu_baseline", d["cpu_baseline"]["value"], "reference arm", r["value"], "e2e ratio %.1f" % (d["e2e"]["value"] / r["value"]), "clocks", d["clocks"])
PY
cat /tmp/rac_only.txt
