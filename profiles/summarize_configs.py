"""gpurun_out/configs_n{1,2,4,8}.jsonl (profiles/configs_bench.py) -> profiles/configs_r02.txt.
Usage: python profiles/summarize_configs.py gpurun_out/configs_n1.jsonl gpurun_out/configs_n2.jsonl ... > profiles/configs_r02.txt"""
import json
import sys

rows = []
for p in sys.argv[1:]:
    with open(p) as f:
        rows += [json.loads(l) for l in f if l.strip()]
cfg = [r for r in rows if r["config"] != "sweep"]
print("BASELINE.json configs on 1/2/4/8 B200 (profiles/configs_bench.py; weak scaling: every rank runs the config's batch; CUDA events, max over ranks)")
print("renders/s and iter/s = serial refinement chain without the CNN; stage = rasterize+concat alone (setup + raster + gt_norm + finalize);")
print("GB/s = algorithmic bytes ((4C+12) W H per image + mesh) / stage time; with-CNN = stock torch CNN fwd+bwd+Adam in the loop (f32, eager, DDP at N>1)")
print()
print("%-4s %2s %5s %9s %12s %10s %9s %8s %10s  %-18s %12s %12s %s" % ("cfg", "N", "B/GPU", "size", "renders/s", "iter/s", "stage us", "GB/s", "Gtri/s", "CNN", "iter/s +CNN", "rend/s +CNN", "CPU oracle renders/s (threads)"))
for name in ("c1", "c2", "c3", "c4"):
    base = None
    for r in sorted([r for r in cfg if r["config"] == name], key=lambda r: r["n_gpus"]):
        if r["n_gpus"] == 1:
            base = r
        cpu = "%.0f (%d)" % (r["cpu_oracle_renders_per_s"], r["cpu_cores"]) if "cpu_oracle_renders_per_s" in r else ""
        sc = " x%.2f" % (r["renders_per_s"] / base["renders_per_s"]) if base and r["n_gpus"] > 1 else ""
        print("%-4s %2d %5d %9s %12.0f %10.1f %9.1f %8.0f %10.2f  %-18s %12s %12s %s" % (
            name, r["n_gpus"], r["B_per_gpu"], "%dx%d" % (r["W"], r["H"]), r["renders_per_s"], r["iters_per_s"], r["stage_us"],
            r["stage_GBps"], r["tri_setup_per_s"] / 1e9, r.get("cnn", "-"),
            "%.2f" % r["iters_per_s_with_cnn"] if "iters_per_s_with_cnn" in r else "-",
            "%.0f" % r["renders_per_s_with_cnn"] if "renders_per_s_with_cnn" in r else "-", cpu + sc))
print()
print("render-only sweep (rac_render_batch: rgb + depth outputs), renders/s by GPU count")
print("%9s %5s %6s %14s %14s %14s %14s %10s" % ("triangles", "px", "B/GPU", "N=1", "N=2", "N=4", "N=8", "Gtri/s N=8"))
sw = [r for r in rows if r["config"] == "sweep"]
keys = sorted({(r["triangles"], r["px"], r["B_per_gpu"]) for r in sw})
for k in keys:
    by = {r["n_gpus"]: r for r in sw if (r["triangles"], r["px"], r["B_per_gpu"]) == k}
    cell = lambda n: "%.0f" % by[n]["renders_per_s"] if n in by else "-"
    print("%9d %5d %6d %14s %14s %14s %14s %10s" % (k[0], k[1], k[2], cell(1), cell(2), cell(4), cell(8),
                                                    "%.1f" % (by[8]["tri_per_s"] / 1e9) if 8 in by else "-"))
