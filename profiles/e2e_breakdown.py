"""Where the end-to-end step (bench.py step_e2e: the reference-facing calls on pinned host buffers) spends its time:
host enqueue time per call (no sync) and the synchronised total.  Diagnostic only.  Usage (GPU box):
python profiles/e2e_breakdown.py [--u8]   (--u8: the gt crop as uint8 codes, rac_gt_norm_u8)"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200"), ROOT]


def main():
    import torch
    import bench
    import rac
    import data_loaders as DL
    import loss as L
    import rotation_representation as RR
    from rac import mesh as M
    from rac import render as R
    R.set_gt_cache(False)
    dev = torch.device("cuda", 0)
    mesh, sets, K = bench.make_inputs(seed=1000)
    B, W, H = bench.B, bench.W, bench.H
    ply = "/tmp/rac_e2e_mesh.ply"
    M.save_ply(ply, mesh)
    paths = [ply] * B
    Ks = torch.tensor(K)
    hs = []
    for s in sets:
        hs.append(dict(gt=(torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8) if "--u8" in sys.argv else torch.rand((B, H, W, 3))).pin_memory(), T=torch.tensor(s["T_init"]).pin_memory(),
                       Tg=torch.tensor(s["T_gt"]).pin_memory(), y=torch.tensor(s["y"]).pin_memory(),
                       pts=torch.tensor(s["pts"]).pin_memory()))
    h_pose = torch.empty((B, 4, 4)).pin_memory()
    h_loss = torch.empty(1).pin_memory()
    names = ["render_batch", "small H2D", "prepare_model_input", "pose_update", "loss", "backward", "D2H+sync"]
    acc = [0.0] * len(names)

    def step(i, rec):
        s = hs[i % len(hs)]
        t = [time.perf_counter()]
        imgs, depths = DL.render_batch_sequential(s["T"].numpy(), paths, Ks, (W, H)); t.append(time.perf_counter())
        T_dev = s["T"].to(dev, non_blocking=True)
        y = s["y"].to(dev, non_blocking=True).requires_grad_(True)
        T_gt = s["Tg"].to(dev, non_blocking=True)
        pts = s["pts"].to(dev, non_blocking=True); t.append(time.perf_counter())
        x = DL.prepare_model_input(imgs, s["gt"], depths, False).to(dev); t.append(time.perf_counter())
        T_new = RR.calculate_T_CO_pred(y, T_dev, bench.ROT, Ks); t.append(time.perf_counter())
        l = L.compute_disentangled_ADD_L1_loss(T_gt, T_new, pts); t.append(time.perf_counter())
        l.backward(); t.append(time.perf_counter())
        h_pose.copy_(T_new.detach(), non_blocking=True)
        h_loss.copy_(l.detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize(); t.append(time.perf_counter())
        if rec:
            for k in range(len(names)):
                acc[k] += t[k + 1] - t[k]
        return x

    for i in range(5):
        step(i, False)
    n = 30
    t0 = time.perf_counter()
    for i in range(n):
        step(i, True)
    tot = (time.perf_counter() - t0) / n
    print("e2e step %.3f ms" % (tot * 1e3))
    for k, nm in enumerate(names):
        print("  %-22s %.3f ms (host time until the call returned)" % (nm, acc[k] / n * 1e3))


def main_driver():
    """The drivers' loop shape and input types (bench.py step_e2e_driver): numpy float64 pageable gt, float64 CPU cam_mats,
    3 iterations per batch, pose back to numpy every iteration."""
    import numpy as np
    import torch
    import bench
    import data_loaders as DL
    import loss as L
    import rotation_representation as RR
    from rac import mesh as M
    dev = torch.device("cuda", 0)
    mesh, sets, K = bench.make_inputs(seed=1000)
    B, W, H = bench.B, bench.W, bench.H
    ply = "/tmp/rac_e2e_mesh.ply"
    M.save_ply(ply, mesh)
    paths = [ply] * B
    Ks = torch.tensor(K.astype(np.float64))
    rng = np.random.default_rng(0)
    hs = [dict(gt=rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8) / 255.0, T=torch.tensor(s["T_init"].astype(np.float64)),
               Tg=torch.tensor(s["T_gt"]), y=torch.tensor(s["y"]).pin_memory(), pts=torch.tensor(s["pts"])) for s in sets]
    names = ["T_gt/pts .to(device)", "render_batch", "torch.tensor(T).to(dev)", "prepare_model_input", "y H2D", "pose_update", "loss",
             "backward", ".cpu().numpy()"]
    acc = [[0.0] * len(names) for _ in range(3)]

    def batch(i, rec):
        s = hs[i % len(hs)]
        T_pred = s["T"]
        t0 = time.perf_counter()
        T_gt = s["Tg"].to(dev); pts = s["pts"].to(dev)
        t_pre = time.perf_counter() - t0
        for j in range(3):
            t = [time.perf_counter()]
            imgs, depths = DL.render_batch(T_pred if j else T_pred.numpy(), paths, Ks, (W, H), False); t.append(time.perf_counter())
            T_dev = torch.tensor(T_pred).to(dev) if j else T_pred.to(dev); t.append(time.perf_counter())
            x = DL.prepare_model_input(imgs, s["gt"], depths, False).to(dev); t.append(time.perf_counter())
            y = s["y"].to(dev, non_blocking=True).requires_grad_(True); t.append(time.perf_counter())
            T_new = RR.calculate_T_CO_pred(y, T_dev, bench.ROT, Ks); t.append(time.perf_counter())
            l = L.compute_disentangled_ADD_L1_loss(T_gt, T_new, pts); t.append(time.perf_counter())
            l.backward(); t.append(time.perf_counter())
            T_pred = T_new.detach().cpu().numpy(); t.append(time.perf_counter())
            if rec:
                acc[j][0] += t_pre if j == 0 else 0.0
                for k in range(len(names) - 1):
                    acc[j][k + 1] += t[k + 1] - t[k]

    for i in range(4):
        batch(i, False)
    n = 16
    t0 = time.perf_counter()
    for i in range(n):
        batch(i, True)
    tot = (time.perf_counter() - t0) / n
    print("e2e batch of 3 iterations %.3f ms (%.3f ms per iteration)" % (tot * 1e3, tot * 1e3 / 3))
    for k, nm in enumerate(names):
        print("  %-26s iter0 %.3f  iter1 %.3f  iter2 %.3f ms (host time until the call returned)" % (nm, acc[0][k] / n * 1e3, acc[1][k] / n * 1e3, acc[2][k] / n * 1e3))


if __name__ == "__main__":
    main_driver() if "--driver" in sys.argv else main()
