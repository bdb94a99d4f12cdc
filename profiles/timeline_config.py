"""Kernel timeline (globaltimer of first CTA start / last CTA end per kernel, diagnostic build -DRAC_PHASE_TIMING) of one
rac_render_concat call at a BASELINE config's shape (profiles/configs_bench.py: c2, c3, c4).  Diagnostic only.
Usage (GPU box): python profiles/build_variants.py timing=-DRAC_PHASE_TIMING; python profiles/timeline_config.py c4"""
import ctypes
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
OUT = os.path.join(PKG, "librac_b200_timing.so")
sys.path[:0] = [PKG, ROOT, os.path.join(ROOT, "profiles")]
NAMES = ["gt_norm (side)", "setup", "raster", "finalize_live"]


def main(name):
    os.environ["RAC_LIB"] = OUT
    import numpy as np
    import torch
    import configs_bench as CB
    import rac
    cfg = CB.configs()[name]
    B, W, H = cfg["B"], cfg["W"], cfg["H"]
    meshes = cfg["meshes"](tempfile.mkdtemp())
    rng = np.random.default_rng(5)
    per_image = [meshes[i] for i in rng.integers(0, len(meshes), B)] if len(meshes) > 1 else meshes[0]
    Ti, _ = CB.poses(B, 0)
    T = torch.tensor(Ti, device="cuda")
    K = torch.tensor(CB.intr(B, W, H), device="cuda")
    gts = [torch.rand((B, H, W, 3), device="cuda") for _ in range(2)]
    outs = [torch.empty((B, 7 if cfg["C7"] else 6, H, W), device="cuda") for _ in range(2)]
    graphs = []
    for i in range(2):
        for _ in range(2):
            rac.render_concat_device(T, per_image, K, (W, H), gts[i], use_norm_depth=cfg["C7"], out=outs[i])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            rac.render_concat_device(T, per_image, K, (W, H), gts[i], use_norm_depth=cfg["C7"], out=outs[i])
        graphs.append(g)
    dbg = ctypes.CDLL(OUT).rac_debug_timeline
    buf = (ctypes.c_ulonglong * 16)()
    acc, n = np.zeros((4, 2)), 0
    for it in range(10):
        torch.cuda.synchronize()
        dbg(None, 1)
        graphs[it % 2].replay()
        torch.cuda.synchronize()
        dbg(buf, 0)
        t = np.array(list(buf), dtype=np.float64).reshape(8, 2)[:4]
        valid = t[:, 1] > 0
        t0 = t[valid, 0].min()
        if it >= 4:
            acc += np.where(valid[:, None], (t - t0) / 1e3, 0.0)
            n += 1
    print("config %s: B=%d %dx%d" % (name, B, W, H))
    for k, nm in enumerate(NAMES):
        print("%-18s start %8.1f us   end %8.1f us   (%.1f us)" % (nm, acc[k, 0] / n, acc[k, 1] / n, (acc[k, 1] - acc[k, 0]) / n))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "c4")
