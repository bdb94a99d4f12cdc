"""Generates tests/golden/raster_golden.npz from the CPU oracle (oracle/raster_ref.c).

These vectors do NOT pin the oracle to the reference (parity of step 1 is UNPINNED: the reference
renders through pyrender/OpenGL, unavailable here and on the GPU box, and ships no golden images).
They pin the oracle to itself, so that (a) a change to raster_ref.c that alters its output is
caught on CPU, and (b) the CUDA kernels are checked on the GPU box against committed bytes.

    python tests/golden/make_raster_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200"), ROOT, os.path.join(ROOT, "tests")]
OUT = os.path.join(HERE, "raster_golden.npz")
W, H = 96, 72


def scenes():
    from rac import mesh as M
    from conftest import intrinsics, sample_poses
    vc = M.icosphere(2)
    rng = np.random.default_rng(0)
    vc.vcolor = np.concatenate([rng.integers(30, 255, (len(vc.verts), 3)), np.full((len(vc.verts), 1), 255)], 1).astype(np.uint8)
    ms = {"bumpy": M.bumpy_sphere(level=3, seed=1), "table": M.table(2), "vcolor": vc,
          "textured": M.with_texture(M.icosphere(3), size=64, seed=3)}
    out = {}
    for i, (name, m) in enumerate(ms.items()):
        T, _ = sample_poses(2, seed=100 + i)
        out[name] = (m, T, intrinsics(2, W, H))
    return out


def main():
    from oracle import raster_oracle as ro
    out = {}
    for name, (m, T, K) in scenes().items():
        rgb, depth, rgba8 = ro.render_batch(m, T, K, (W, H))
        out[f"{name}_T"], out[f"{name}_K"] = T, K
        out[f"{name}_rgba8"], out[f"{name}_depth"] = rgba8, depth
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
