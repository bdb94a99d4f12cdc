"""A/B of the rasterize+concat stage under the library's tuning knobs (RAC_FOLD, RAC_WAVES, RAC_WAVE_CTAS,
RAC_GT_CLUSTERS, ...) at the bench workload: every configuration runs in its own process (knobs are read once),
times the stage as a CUDA-graph replay (4 graphs in rotation, like bench.py) and prints a checksum of x so that the
variants can be seen to agree bit for bit.  Diagnostic only.
Usage (GPU box): python profiles/ab_stage.py [--depth] "RAC_FOLD=0" "RAC_FOLD=1 RAC_WAVES=2" ...
"""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
sys.path[:0] = [PKG, ROOT]


def child(use_depth):
    import torch
    import bench
    import rac
    mesh, sets, K = bench.make_inputs(seed=1000)
    B, W, H = bench.B, bench.W, bench.H
    g0 = torch.Generator(device="cuda").manual_seed(7)
    gts = [torch.rand((B, H, W, 3), device="cuda", generator=g0) for _ in range(4)]
    outs = [torch.empty((B, 7 if use_depth else 6, H, W), device="cuda") for _ in range(4)]
    Ts = [torch.tensor(s["T_init"], device="cuda") for s in sets]
    Kd = torch.tensor(K, device="cuda")
    graphs = []
    for i in range(4):
        for _ in range(2):
            rac.render_concat_device(Ts[i], mesh, Kd, (W, H), gts[i], use_norm_depth=use_depth, out=outs[i])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            rac.render_concat_device(Ts[i], mesh, Kd, (W, H), gts[i], use_norm_depth=use_depth, out=outs[i])
        graphs.append(g)
    for i in range(12):
        graphs[i % 4].replay()
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 40
        t0.record()
        for i in range(n):
            graphs[i % 4].replay()
        t1.record()
        torch.cuda.synchronize()
        best = min(best, t0.elapsed_time(t1) * 1e3 / n)
    h = hashlib.sha1()
    for o in outs:
        h.update(o.cpu().numpy().tobytes())
    print("%8.1f us   x sha1 %s" % (best, h.hexdigest()[:12]))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        child(sys.argv[2] == "1")
    else:
        args = sys.argv[1:]
        use_depth = "--depth" in args
        args = [a for a in args if a != "--depth"] or [""]
        for cfg in args:
            env = dict(os.environ)
            for kv in cfg.split():
                k, v = kv.split("=")
                env[k] = v
            r = subprocess.run([sys.executable, __file__, "--child", "1" if use_depth else "0"], capture_output=True, text=True, env=env)
            print("%-58s %s" % (cfg or "(defaults)", (r.stdout.strip().splitlines() or [r.stderr[-400:]])[-1]), flush=True)
