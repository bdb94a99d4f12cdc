"""Kernel timeline of one rac_render_concat call at the bench workload (diagnostic build): for each kernel the
globaltimer of its first CTA start and last CTA end, relative to the earliest start, for graph replays.
Shows how much of the forked gt/background work overlaps setup + raster.  Diagnostic only.
Usage (GPU box): python profiles/kernel_timeline.py [RAC_DEBUG_SKIP mask]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
OUT = os.environ.get("RAC_TIMING_LIB", os.path.join(PKG, "librac_b200_timing.so"))
sys.path[:0] = [PKG, ROOT]
NAMES = ["gt_norm (side)", "setup", "raster", "finalize_live"]


def main():
    os.environ["RAC_LIB"] = OUT
    if len(sys.argv) > 1:
        os.environ["RAC_DEBUG_SKIP"] = sys.argv[1]
    import numpy as np
    import torch
    import bench
    import rac
    from rac import _lib
    L = _lib.lib()
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
    dbg = ctypes.CDLL(OUT).rac_debug_timeline      # same handle: dlopen of an already loaded library
    buf = (ctypes.c_ulonglong * 16)()
    acc = np.zeros((4, 2))
    n = 0
    cta_start, cta_end = [], []
    from rac import render as R
    cyc = 0.0
    for it in range(12):
        torch.cuda.synchronize()
        dbg(None, 1)
        graphs[it % 4].replay()
        torch.cuda.synchronize()
        dbg(buf, 0)
        late = buf[8]
        if it >= 4:
            g = buf[4 * 1 + 0]      # raster first start (index 2*2)
            ncta = 592
            ep = buf[15]
            cta_start.append(buf[9] / ncta - (buf[4] - ep)); cta_end.append((buf[5] - ep) - buf[10] / ncta)
        t = np.array(list(buf), dtype=np.float64).reshape(8, 2)[:4]
        valid = t[:, 1] > 0
        t0 = t[valid, 0].min()
        if it >= 4:
            acc += np.where(valid[:, None], (t - t0) / 1e3, 0.0)
            n += 1
            ws = max(R._ws_cache.values(), key=lambda w: w.numel())
            base = (ws.data_ptr() + 255) // 256 * 256 - ws.data_ptr()
            off = base + (64 * B + 255) // 256 * 256            # WorkCtl follows the ImgStats block
            ctl = ws[off:off + 256].cpu().numpy().view(np.uint32)
            cyc += float(ctl[2:64].copy().view(np.uint64)[:13].sum())
    print("raster: mean CTA start lag %.1f us after the first CTA, mean CTA finish %.1f us before the last (592 CTAs assumed)"
          % (np.mean(cta_start) / 1e3, np.mean(cta_end) / 1e3))
    print("raster CTAs that started > 10 us after the first (last replay): %d" % late)
    print("raster CTA cycles (sum over CTAs, thread 0 clock64): %.1f M per call" % (cyc / n / 1e6))
    for nm, (a, b) in zip(NAMES, acc / n):
        print("%-18s start %7.1f us   end %7.1f us   (%.1f us)" % (nm, a, b, b - a))


if __name__ == "__main__":
    main()
