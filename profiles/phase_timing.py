"""Phase timeline of raster_kernel: builds librac_b200_timing.so with -DRAC_PHASE_TIMING (thread 0 of every CTA
accumulates clock64() deltas per phase), runs the bench workload's render a few times and prints the share of
CTA time per phase.  Diagnostic only.  Usage (GPU box): python profiles/phase_timing.py [--build-only]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
OUT = os.environ.get("RAC_TIMING_LIB", os.path.join(PKG, "librac_b200_timing.so"))
sys.path[:0] = [PKG, ROOT]

NAMES = ["item decode", "descriptors+scan+gather", "coverage (flat)", "  barrier after flat", "medium tris", "large list",
         "R1 sample analysis", "  barrier", "R2 shading", "  barrier", "R3 resolve+store", "reduce+atomics", "  next-item hand-over"]

def main():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_rac_build", os.path.join(PKG, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if not os.path.exists(OUT) or "--build" in sys.argv or "--build-only" in sys.argv:
        mod.build_variant(OUT, ["-DRAC_PHASE_TIMING"])
    if "--build-only" in sys.argv:
        return
    os.environ["RAC_LIB"] = OUT
    if len(sys.argv) > 1 and sys.argv[1].isdigit():
        os.environ["RAC_DEBUG_SKIP"] = sys.argv[1]
    import numpy as np
    import torch
    import bench
    import rac
    from rac import render as R
    mesh, sets, K = bench.make_inputs(seed=1000)
    B, W, H = bench.B, bench.W, bench.H
    gt = torch.rand((B, H, W, 3), device="cuda")
    tot = np.zeros(31, dtype=np.float64)
    cls_n = np.zeros(16, dtype=np.float64)
    for it in range(6):
        rac.render_concat_device(sets[it % len(sets)]["T_init"], mesh, K, (W, H), gt)
        torch.cuda.synchronize()
        ws = max(R._ws_cache.values(), key=lambda w: w.numel())
        base = (ws.data_ptr() + 255) // 256 * 256 - ws.data_ptr()
        off = base + (64 * B + 255) // 256 * 256            # WorkCtl follows the ImgStats block
        ctl = ws[off:off + 512].cpu().numpy().view(np.uint32)
        pad = ctl[2:64].copy().view(np.uint64)
        if it >= 2:
            tot += pad.astype(np.float64)
        if it >= 2:
            cls_n += ctl[64:80].astype(np.float64)
        print("n_work", int(ctl[64:80].sum()), "by load class (heaviest first):", ctl[64:80].tolist())
    s = tot[:len(NAMES)].sum()
    for n, v in zip(NAMES, tot):
        print("%-28s %6.2f%%  %10.0f cycles/iter" % (n, 100 * v / s, v / 4))
    return tot, cls_n


def per_class(tot, cls_n):
    print("bin duration by load class (class k: load in [32 (15-k), 32 (16-k)) refs; thread-0 cycles per bin, 4 CTAs/SM co-resident)")
    for k in range(16):
        if cls_n[k] > 0:
            print("  class %2d  bins/iter %6.1f   cycles/bin %8.0f" % (k, cls_n[k] / 4, tot[14 + k] / cls_n[k]))


if __name__ == "__main__":
    res = main()
    if res is not None:      # (--build-only returns nothing)
        per_class(*res)
