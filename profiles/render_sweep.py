"""Render-only sweep (BASELINE.json configs[4]): triangles x poses per batch x resolution on one
B200, RGBA8 + depth32 out (algorithmic 8*W*H bytes per image), poses from rac_sample_poses, device-timed with CUDA
events (3 warm-up calls, 5 timed).  The CPU column is the oracle (oracle/raster_ref.c, all host threads) on a bounded
sample of <= 8 poses of the same batch.  Workspace is capped at 40 GB per case.  Diagnostic / evidence only.
Usage (GPU box): python profiles/render_sweep.py [--quick] > profiles/render_sweep_r01.txt"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200"), ROOT]

CFG = dict(distance_cam_to_world=2.5, distance_cam_to_world_deviation=0.1, world_to_object_gt_transl_deviation=0.1,
           world_to_object_transl_deviation=0.1, world_to_object_angle_deviation=30)


def main():
    import numpy as np
    import torch
    import rac
    from rac import mesh as M
    from oracle import raster_oracle
    quick = "--quick" in sys.argv
    L = rac.lib()
    dev = torch.device("cuda", 0)
    meshes = {}

    def get_mesh(T):
        if T not in meshes:
            level = 3
            while 20 * 4 ** level < T:
                level += 1
            meshes[T] = M.bumpy_sphere(level=level, seed=1, decimate_to=T)
        return meshes[T]

    tri_list = [1000, 5000, 20000, 100000]
    pose_list = [1, 64, 512, 4096]
    cases = [(T, B, 320, 240) for T in tri_list for B in pose_list]
    cases += [(10000, 64, S, S) for S in (224, 320, 512, 640)]
    if quick:
        cases = [(1000, 64, 320, 240), (20000, 512, 320, 240), (10000, 64, 640, 640)]
    g = torch.Generator(device=dev)
    g.manual_seed(5000)
    print("%8s %6s %9s %10s %12s %12s %10s %12s" % ("tris", "poses", "WxH", "ms/call", "renders/s", "Mtri/s", "GB/s(8WH)", "cpu rend/s"))
    for T, B, W, H in cases:
        mesh = get_mesh(T)
        nT = len(mesh.faces)
        need = L.rac_render_workspace_bytes(B, W, H, nT)
        if need > 40e9:
            print("%8d %6d %9s   skipped: workspace %.1f GB > 40 GB cap" % (nT, B, "%dx%d" % (W, H), need / 1e9))
            continue
        T_init, _ = rac.sample_poses_device(B, CFG, generator=g)
        fx = 50.0 / (36.0 / W)
        K = torch.tensor([[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]], dtype=torch.float32, device=dev).repeat(B, 1, 1)
        out = None
        for _ in range(3):
            out = rac.render_batch_device(T_init, mesh, K, (W, H), want_rgb=False, want_depth=True, want_rgba8=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            out = rac.render_batch_device(T_init, mesh, K, (W, H), want_rgb=False, want_depth=True, want_rgba8=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        cov = float((out["depth"] > 0).float().mean())
        assert 0.01 < cov < 0.9, cov
        nb = min(B, 8 if nT <= 20000 else 2)
        Th, Kh = T_init[:nb].cpu().numpy(), K[:nb].cpu().numpy()
        t0 = time.perf_counter()
        _, d_o, _ = raster_oracle.render_batch(mesh, Th, Kh, (W, H))
        cpu = nb / (time.perf_counter() - t0)
        mism = float(((d_o > 0) != (out["depth"][:nb].cpu().numpy() > 0)).mean())
        assert mism <= 1e-3, mism
        print("%8d %6d %9s %10.3f %12.0f %12.1f %10.1f %12.1f" % (nT, B, "%dx%d" % (W, H), ms, B / ms * 1e3, B * nT / ms / 1e3,
                                                             8.0 * W * H * B / ms / 1e6, cpu), flush=True)
        del out
        torch.cuda.empty_cache()
        from rac import render as R
        R._ws_cache.clear()
    print("cpu column: oracle/raster_ref.c with %d OpenMP threads on <= 8 poses of the batch; coverage of the sampled poses checked against it (<= 0.1%% of pixels differ)"
          % raster_oracle.num_threads())


if __name__ == "__main__":
    main()
