"""BASELINE.json configs[0..4] timed on N GPUs (one process per GPU; launch with torch.distributed.run for N > 1):

  c1  configs/example_config.py shape: B=8, 320x320, one untextured ~10k-triangle mesh, 6D, scaled disentangled loss,
      normalised depth plane (C=7), 1 refinement iteration, BaseNet-shaped CNN
  c2  old-configs/effnet_b3_6d.py shape at the metric's size: B=64, 320x240, untextured ~10k triangles, 6D, disentangled
      loss, 3 iterations, EfficientNet-B3 (torchvision, 6-channel stem)
  c3  old-configs/effnetv2_m_svd.py: B=128, 320x320, SVD head, ADD-L1, TEXTURIZED meshes (multi-part .glb scenes read by
      rac/gltf.py), EfficientNetV2-M (torchvision), DDP at N > 1
  c4  old-configs/effnetv2_m_dloss2.py: B=256, 320x320, SVD head, disentangled loss, 5 iterations, MIXED meshes (10
      distinct procedural meshes of 2k-40k triangles, drawn per sample)
  sweep  render-only grid: {1k, 10k, 100k} triangles x {1, 64, 1024, 4096} poses x {224, 320, 640} px

Per config, weak scaling (every rank runs the config's batch): renders/s of the SERIAL refinement chain (render ->
pose update + loss + backward -> T_pred <- T_new, one graph replay per iteration, a fresh batch every `iters`
iterations), the rasterize+concat stage time on its own, its algorithmic GB/s (36 or 40 * W * H per image + mesh
bytes), the triangle setup rate, iterations/s with the stock CNN in the loop (forward + backward + Adam, DDP gradient
all-reduce over NCCL at N > 1; the CNN runs in micro-batches of 64 so that activations stay bounded), and — on rank 0 at
N = 1 — the CPU oracle's renders/s on a bounded sample with its thread count.
Times are CUDA events on the launching stream, max over ranks.  Diagnostic / evidence script: its table goes to
profiles/configs_r02.txt.

Usage: python profiles/configs_bench.py [--configs c1,c2,c3,c4,sweep] [--no-cnn] [--no-cpu]
       python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P profiles/configs_bench.py ...
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
sys.path[:0] = [PKG, ROOT, os.path.join(ROOT, "tests")]

import numpy as np  # noqa: E402

SCENE_CFG = dict(distance_cam_to_world=2.5, distance_cam_to_world_deviation=0.1, world_to_object_gt_transl_deviation=0.1,
                 world_to_object_transl_deviation=0.1, world_to_object_angle_deviation=30)


def poses(B, seed):
    import se3_helpers
    np.random.seed(seed)
    ti, tg = [], []
    for _ in range(B):
        a, b = se3_helpers.get_T_CO_init_and_gt(SCENE_CFG)
        ti.append(a.data[0]); tg.append(b.data[0])
    return np.array(ti, np.float32), np.array(tg, np.float32)


def intr(B, W, H):
    fx = 50.0 / (36.0 / W)
    return np.tile(np.array([[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]], np.float32), (B, 1, 1))


def level_for(tris):
    return 3 if tris <= 1280 else 4 if tris <= 5120 else 5 if tris <= 20480 else 6 if tris <= 81920 else 7


def glb_scene(tmpdir, k, tris):
    """A 'texturized' object: textured bumpy body + two factor-coloured boxes, written as a multi-primitive .glb."""
    from rac import gltf
    from rac import mesh as M
    body = M.with_texture(M.bumpy_sphere(level=level_for(tris), seed=20 + k, decimate_to=tris - 200), size=256, seed=k)
    b1, b2 = M.box(0.5, 0.12, 0.12, 2), M.box(0.12, 0.5, 0.12, 2)
    path = os.path.join(tmpdir, "texturized_%02d.glb" % k)
    if os.environ.get("RAC_C3_NOTEX"):          # diagnostic: the same geometry without the texture (cost of R8)
        body.tex = None
    gltf.save_glb(path, [dict(verts=body.verts, faces=body.faces, uv=body.uv, tex=body.tex, material=(1, 1, 1, 1, 0.0, 0.6)),
                         dict(verts=b1.verts + np.float32([0, 0, 0.42]), faces=b1.faces, material=(0.7, 0.3, 0.2, 1, 0.3, 0.5)),
                         dict(verts=b2.verts + np.float32([0, 0, -0.42]), faces=b2.faces, material=(0.2, 0.4, 0.7, 1, 0.8, 0.4))])
    return path


def make_cnn(name, C, D, H, W, dev):
    import torch.nn as nn
    if name == "basenet":
        def block(ci, co):
            return nn.Sequential(nn.Conv2d(ci, co, 3, 1, 1), nn.BatchNorm2d(co), nn.ReLU(),
                                 nn.Conv2d(co, co, 3, 1, 1), nn.BatchNorm2d(co), nn.ReLU(), nn.MaxPool2d(2, 2))
        ch = [C, 32, 48, 64, 64, 32]
        return nn.Sequential(*[block(ch[i], ch[i + 1]) for i in range(5)], nn.Flatten(), nn.Linear(32 * (H // 32) * (W // 32), D)).to(dev)
    import torchvision
    net = (torchvision.models.efficientnet_b3 if name == "effnet_b3" else torchvision.models.efficientnet_v2_m)(weights=None, num_classes=D)
    stem = net.features[0][0]
    net.features[0][0] = nn.Conv2d(C, stem.out_channels, stem.kernel_size, stem.stride, stem.padding, bias=False)
    return net.to(dev)


def run_config(name, cfg, args, dev, world, rank, tmpdir):
    import torch
    import torch.distributed as dist
    import rac
    from oracle import raster_oracle as ro
    B, W, H, rot, loss, C7, iters = cfg["B"], cfg["W"], cfg["H"], cfg["rot"], cfg["loss"], cfg["C7"], cfg["iters"]
    meshes = cfg["meshes"](tmpdir)
    rng = np.random.default_rng(5 + rank)
    per_image = [meshes[i] for i in rng.integers(0, len(meshes), B)] if len(meshes) > 1 else meshes[0]
    D = {"6D": 9, "SVD": 12}[rot]
    NS = 2
    steps = []
    tri_total = 0
    for s in range(NS):
        st = rac.RenderCompareStep(per_image, B, (W, H), rot_repr=rot, loss=loss, n_points=1000, use_norm_depth=C7)
        Ti, Tg = poses(B, 100 * rank + s)
        y = (0.05 * rng.normal(size=(B, D))).astype(np.float32)
        y[:, 0] += 1; y[:, 4] += 1
        if rot == "SVD":
            y[:, 8] += 1
        y[:, -3:-1] *= 0.2; y[:, -1] = 1.0 + 0.02 * rng.normal(size=B)
        st.T_pred.copy_(torch.tensor(Ti)); st.T_prev.copy_(torch.tensor(Ti)); st.K.copy_(torch.tensor(intr(B, W, H)))
        st.gt.copy_(torch.randint(0, 256, (B, H, W, 3), device=dev).float() / 255.0)
        st.y.copy_(torch.tensor(y)); st.T_gt.copy_(torch.tensor(Tg))
        st.pts.copy_(torch.tensor(rng.normal(size=(B, 1000, 3)).astype(np.float32) * 0.3))
        st.capture()
        st.T_init = torch.tensor(Ti, device=dev)
        steps.append(st)
    from rac.mesh import device_mesh
    ms = per_image if isinstance(per_image, list) else [per_image] * B
    tri_total = sum(device_mesh(m).T for m in ms)
    vert_total = sum(device_mesh(m).V for m in set(ms))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0.record()
        for i in range(n):
            fn(i)
        t1.record()
        barrier()
        t = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def chain(i):
        st = steps[i % NS]
        if (i // NS) % iters == 0:
            st.T_pred.copy_(st.T_init); st.T_prev.copy_(st.T_init)
        st.run_chain()

    n = max(12, min(60, int(6000 / max(B, 1))))
    for i in range(4):
        chain(i)
    chain_ms = timed(chain, n) / n
    for i in range(3):
        steps[i % NS].run_render()
    stage_ms = timed(lambda i: steps[i % NS].run_render(), n) / n
    Cc = 7 if C7 else 6
    algo = (4.0 * Cc + 12.0) * W * H * B + 12.0 * (vert_total + tri_total / B * 1.0) + 100.0 * B
    row = {"config": name, "n_gpus": world, "B_per_gpu": B, "W": W, "H": H, "triangles_per_batch": int(tri_total),
           "renders_per_s": world * B / (chain_ms * 1e-3), "iters_per_s": world / (chain_ms * 1e-3), "chain_ms": chain_ms,
           "stage_us": stage_ms * 1e3, "stage_GBps": algo / (stage_ms * 1e-3) / 1e9,
           "tri_setup_per_s": tri_total / (stage_ms * 1e-3)}
    # ---- stock CNN in the loop (DDP at N > 1) ----
    if not args.no_cnn and cfg.get("cnn"):
        torch.manual_seed(0)
        net = make_cnn(cfg["cnn"], Cc, D, H, W, dev)
        if world > 1:
            net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[dev.index], broadcast_buffers=False)   # (several micro-batch forwards per backward)
        opt = torch.optim.Adam(net.parameters(), lr=3e-4)
        MB = 64
        idm = torch.zeros(D, device=dev)
        idm[0] = 1; idm[4] = 1
        if rot == "SVD":
            idm[8] = 1

        def with_cnn(i):
            st = steps[i % NS]
            if (i // NS) % iters == 0:
                st.T_pred.copy_(st.T_init); st.T_prev.copy_(st.T_init)
            opt.zero_grad(set_to_none=True)
            st.run_render()
            outs = [net(st.x[b0:b0 + MB]) for b0 in range(0, B, MB)] if B > MB else [net(st.x)]
            out = torch.cat(outs, 0)
            y = torch.cat([out[:, :-3] * 0.02, out[:, -3:-1] * 0.01, 1.0 + 0.01 * torch.tanh(out[:, -1:])], 1) + idm
            st.y.copy_(y.detach())
            st.run_pose_loss()
            y.backward(st.grad_y)
            opt.step()
            st.T_pred.copy_(st.T_new); st.T_prev.copy_(st.T_new)

        nc = 6 if B >= 128 else 10
        for i in range(2):
            with_cnn(i)
        cnn_ms = timed(with_cnn, nc) / nc
        row.update({"cnn": cfg["cnn"] + (" + DDP" if world > 1 else ""), "iters_per_s_with_cnn": world / (cnn_ms * 1e-3),
                    "renders_per_s_with_cnn": world * B / (cnn_ms * 1e-3), "with_cnn_ms": cnn_ms})
        del net, opt
        torch.cuda.empty_cache()
    # ---- CPU oracle column (rank 0, N = 1): bounded sample of the same render ----
    if world == 1 and not args.no_cpu:
        nb = min(B, 8)
        from rac.mesh import host_mesh
        ms_o = [host_mesh(m) if isinstance(m, str) else m for m in ms[:nb]]
        Ti, _ = poses(nb, 1)
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 4.0:
            ro.render_batch_mixed(ms_o, Ti, intr(nb, W, H), (W, H))
            reps += 1
        dt = time.perf_counter() - t0
        row.update({"cpu_oracle_renders_per_s": nb * reps / dt, "cpu_cores": ro.num_threads(), "cpu_sample": "%d x %d poses" % (reps, nb)})
    for st in steps:
        del st
    torch.cuda.empty_cache()
    return row


def run_sweep(args, dev, world, rank):
    import torch
    import torch.distributed as dist
    import rac
    from rac import mesh as M
    rows = []
    for tris in (1000, 10000, 100000):
        mesh = M.bumpy_sphere(level=level_for(tris), seed=3, decimate_to=tris)
        for W in (224, 320, 640):
            for B in (1, 64, 1024, 4096):
                if B * W * W * 16 > 40e9:          # keep the outputs + workspace well inside HBM
                    continue
                Ti, _ = poses(min(B, 256), 7 + rank)
                T = torch.tensor(np.tile(Ti, ((B + len(Ti) - 1) // len(Ti), 1, 1))[:B], device=dev)
                K = torch.tensor(intr(B, W, W), device=dev)
                f = lambda: rac.render_batch_device(T, mesh, K, W, want_depth=True, want_rgb=True)
                for _ in range(2):
                    f()
                n = 20 if B <= 64 else (6 if B <= 1024 else 3)
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                t0.record()
                for _ in range(n):
                    f()
                t1.record()
                torch.cuda.synchronize()
                t = torch.tensor([t0.elapsed_time(t1) / n], device=dev, dtype=torch.float64)
                if world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t[0])
                rows.append({"config": "sweep", "n_gpus": world, "triangles": len(mesh.faces), "px": W, "B_per_gpu": B,
                             "renders_per_s": world * B / (ms * 1e-3), "tri_per_s": world * B * len(mesh.faces) / (ms * 1e-3), "ms": ms})
                del T, K
                torch.cuda.empty_cache()
    return rows


def configs():
    from rac import mesh as M
    one10k = lambda tmp: [M.bumpy_sphere(level=5, seed=1, decimate_to=10000)]
    glbs = lambda tmp: [glb_scene(tmp, k, 12000 + 1500 * k) for k in range(6)]
    mixed = lambda tmp: [M.bumpy_sphere(level=level_for(t), seed=30 + i, decimate_to=t)
                         for i, t in enumerate((2000, 4000, 6000, 8000, 10000, 14000, 18000, 24000, 32000, 40000))]
    return {
        "c1": dict(B=8, W=320, H=320, rot="6D", loss="add_l1_disentl_scaled", C7=True, iters=1, meshes=one10k, cnn="basenet"),
        "c2": dict(B=64, W=320, H=240, rot="6D", loss="add_l1_disentangled", C7=False, iters=3, meshes=one10k, cnn="effnet_b3"),
        "c3": dict(B=128, W=320, H=320, rot="SVD", loss="add_l1", C7=False, iters=3, meshes=glbs, cnn="effnetv2_m"),
        "c4": dict(B=256, W=320, H=320, rot="SVD", loss="add_l1_disentangled", C7=False, iters=5, meshes=mixed, cnn="effnetv2_m"),
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c2,c3,c4,sweep")
    ap.add_argument("--no-cnn", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import tempfile
    import torch
    import torch.distributed as dist
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    tmpdir = tempfile.mkdtemp(prefix="rac_cfg_r%d_" % rank)
    rows = []
    cfgs = configs()
    for name in args.configs.split(","):
        new = run_sweep(args, dev, world, rank) if name == "sweep" else [run_config(name, cfgs[name], args, dev, world, rank, tmpdir)]
        rows += new
        if rank == 0:
            for r in new:
                print(json.dumps(r), flush=True)
    if rank == 0 and args.out:
        with open(args.out, "a") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
