"""Stage breakdown of rac_render_concat at the bench workload: CUDA-event time of the eager call with individual
kernels skipped (diagnostic build, -DRAC_PHASE_TIMING enables the RAC_DEBUG_SKIP mask).  Diagnostic only.
Usage (GPU box): python profiles/stage_breakdown.py"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
OUT = os.path.join(PKG, "librac_b200_timing.so")
sys.path[:0] = [PKG, ROOT]

CASES = [(0, "all (product order: gt kernel forked after setup)"), (32, "all, gt kernel forked BEFORE setup"), (64, "all, gt kernel one cluster per image"),
         (16, "side work on the main stream (no fork)"),
         (1, "no side work"), (8, "no finalize"), (9, "no side work, no finalize"),
         (13 + 16, "setup only"), (14 + 16, "side work only"), (14 + 16 + 64, "side work only, cluster per image"), (15 + 16, "memset only")]


def child(mask):
    os.environ["RAC_LIB"] = OUT
    os.environ["RAC_DEBUG_SKIP"] = str(mask)
    import torch
    import bench
    import rac
    mesh, sets, K = bench.make_inputs(seed=1000)
    B, W, H = bench.B, bench.W, bench.H
    gts = [torch.rand((B, H, W, 3), device="cuda") for _ in range(4)]
    outs = [torch.empty((B, 6, H, W), device="cuda") for _ in range(4)]
    Ts = [torch.tensor(s["T_init"], device="cuda") for s in sets]
    Kd = torch.tensor(K, device="cuda")
    graphs = []
    for i in range(4):
        for _ in range(2):
            rac.render_concat_device(Ts[i], mesh, Kd, (W, H), gts[i], out=outs[i])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            rac.render_concat_device(Ts[i], mesh, Kd, (W, H), gts[i], out=outs[i])
        graphs.append(g)
    for i in range(8):
        graphs[i % 4].replay()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    n = 40
    t0.record()
    for i in range(n):
        graphs[i % 4].replay()
    t1.record()
    torch.cuda.synchronize()
    print("%8.1f us" % (t0.elapsed_time(t1) * 1e3 / n))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        child(int(sys.argv[1]))
    else:
        for mask, name in CASES:
            r = subprocess.run([sys.executable, __file__, str(mask)], capture_output=True, text=True)
            print("%-36s %s" % (name, (r.stdout.strip().splitlines() or [r.stderr[-300:]])[-1]), flush=True)
