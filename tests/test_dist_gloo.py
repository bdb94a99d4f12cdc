"""CPU, world_size 2 over gloo: the host-side multi-GPU logic (rac.dist) — contiguous batch sharding, the
global loss mean with its gradient, and the Q2 first-sample-intrinsics broadcast.  The per-shard pose/loss math is
the CPU oracle here (the CUDA kernels are covered by the -m gpu tests); what is checked is that sharded == unsharded."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _inputs(B=7, N=50):
    g = torch.Generator().manual_seed(0)
    y = 0.1 * torch.randn((B, 9), generator=g)
    y[:, 0] += 1; y[:, 4] += 1; y[:, 8] = 1.0
    T = torch.eye(4).repeat(B, 1, 1)
    T[:, :3, 3] = torch.tensor([0.05, -0.02, 2.5]) + 0.05 * torch.randn((B, 3), generator=g)
    Tg = T.clone()
    Tg[:, 0, 3] += 0.03
    pts = 0.3 * torch.randn((B, N, 3), generator=g)
    Ks = torch.tensor([[400.0, 0, 160], [0, 410.0, 120], [0, 0, 1]]).repeat(B, 1, 1)
    Ks[1:, 0, 0] = 999.0          # only the GLOBAL first sample's intrinsics may be used (Q2)
    return y, T, Tg, pts, Ks


def _worker(rank, ws, port, out_dir):
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    from rac import dist as rd
    from oracle import pose_loss_ref as R
    y, T, Tg, pts, Ks = _inputs()
    B = y.shape[0]
    lo, hi = rd.shard_bounds(B)
    ys = y[lo:hi].clone().requires_grad_(True)
    fx, fy = rd.first_sample_intrinsics(Ks[lo:hi])
    Kuse = torch.tensor([[fx, 0, 0], [0, fy, 0], [0, 0, 1]]).repeat(hi - lo, 1, 1)
    Tn = R.pose_update(ys, T[lo:hi], "6D", Kuse)
    per = R.add_l1(Tg[lo:hi], Tn, pts[lo:hi], batch_mean=False)
    loss = rd.global_loss_mean(per)
    loss.backward()
    shard = rd.shard({"y": y, "paths": ["m%d" % i for i in range(B)]})
    assert len(shard["paths"]) == hi - lo and shard["paths"][0] == "m%d" % lo
    np.savez(os.path.join(out_dir, "r%d.npz" % rank), lo=lo, hi=hi, loss=loss.item(), grad=ys.grad.numpy(), fx=fx, fy=fy)
    dist.destroy_process_group()


def test_shard_bounds_cover_the_batch():
    sys.path.insert(0, PKG)
    from rac import dist as rd
    for B in (0, 1, 7, 64, 65):
        for ws in (1, 2, 3, 8):
            cuts = [rd.shard_bounds(B, r, ws) for r in range(ws)]
            assert cuts[0][0] == 0 and cuts[-1][1] == B
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(ws - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1


@pytest.mark.timeout(120)
def test_sharded_loss_and_grads_equal_unsharded(tmp_path):
    ws = 2
    mp.spawn(_worker, args=(ws, _free_port(), str(tmp_path)), nprocs=ws, join=True)
    sys.path[:0] = [PKG, ROOT]
    from oracle import pose_loss_ref as R
    y, T, Tg, pts, Ks = _inputs()
    yf = y.clone().requires_grad_(True)
    full = R.add_l1(Tg, R.pose_update(yf, T, "6D", Ks), pts)          # reference semantics: Ks[0] for everyone
    full.backward()
    parts = [np.load(os.path.join(str(tmp_path), "r%d.npz" % r)) for r in range(ws)]
    assert parts[0]["lo"] == 0 and parts[-1]["hi"] == y.shape[0] and parts[0]["hi"] == parts[1]["lo"]
    for p in parts:
        assert p["fx"] == 400.0 and p["fy"] == 410.0
        np.testing.assert_allclose(p["loss"], full.item(), rtol=1e-6)
    np.testing.assert_allclose(np.concatenate([p["grad"] for p in parts]), yf.grad.numpy(), rtol=1e-5, atol=1e-8)
