"""Generates tests/golden/pose_loss_golden.npz by EXECUTING THE REFERENCE ITSELF
(/root/reference/rotation_representation.py and loss.py, imported unmodified, CPU torch).
Run in the build container only (the reference does not travel to the GPU box):

    python tests/golden/make_pose_loss_golden.py
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pose_loss_golden.npz")


def load_reference():
    saved_path = list(sys.path)
    saved_mods = {k: sys.modules.pop(k) for k in ("rotation_representation", "loss") if k in sys.modules}
    sys.path.insert(0, REF)
    try:
        import rotation_representation as rr
        import loss as ls
        assert rr.__file__.startswith(REF) and ls.__file__.startswith(REF)
    finally:
        sys.path[:] = saved_path
        for k in ("rotation_representation", "loss"):
            sys.modules.pop(k, None)
        sys.modules.update(saved_mods)
    return rr, ls


def random_pose(g, B):
    q = torch.randn(B, 4, generator=g)
    q = q / q.norm(dim=1, keepdim=True)
    w, x, y, z = q.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                     2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                     2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], 1).reshape(B, 3, 3)
    T = torch.eye(4).repeat(B, 1, 1)
    T[:, :3, :3] = R
    T[:, :3, 3] = torch.randn(B, 3, generator=g) * 0.1 + torch.tensor([0.0, 0.0, 2.5])
    return T


def main():
    rr, ls = load_reference()
    import contextlib, io
    g = torch.Generator().manual_seed(1234)
    B, N = 8, 50
    out = {}
    Ks = torch.tensor([[444.44, 0, 160.0], [0, 444.44, 160.0], [0, 0, 1]]).repeat(B, 1, 1).double()
    T_in = random_pose(g, B)
    T_gt = random_pose(g, B)
    T_prev = random_pose(g, B)
    pts = torch.randn(B, N, 3, generator=g) * 0.3
    out.update(Ks=Ks.numpy(), T_in=T_in.numpy(), T_gt=T_gt.numpy(), T_prev=T_prev.numpy(), pts=pts.numpy())
    for name, D in (("6D", 9), ("SVD", 12), ("3D", 6)):
        y = (0.1 * torch.randn(B, D, generator=g))
        if name == "6D":
            y[:, 0] += 1.0; y[:, 4] += 1.0
        if name == "SVD":
            y[:, 0] += 1.0; y[:, 4] += 1.0; y[:, 8] += 1.0
        y[:, D - 1] = 1.0 + 0.05 * torch.randn(B, generator=g)
        if name == "SVD":   # one sample with det < 0 exercises the reflection branch
            y[3, :9] = torch.tensor([1.0, 0.1, 0, 0.2, 1.0, 0, 0, 0.1, -0.7])
        y.requires_grad_(True)
        out[f"y_{name}"] = y.detach().numpy().copy()
        for tag, Tin in (("f32", T_in), ("f64", T_in.double())):
            T_new = rr.calculate_T_CO_pred(y, Tin, name, Ks)
            out[f"Tnew_{name}_{tag}"] = T_new.detach().numpy().copy()
        T_new = rr.calculate_T_CO_pred(y, T_in, name, Ks)
        # gradient of a fixed random linear functional of T_new
        Wt = torch.randn(B, 4, 4, generator=g)
        out[f"W_{name}"] = Wt.numpy()
        (gy,) = torch.autograd.grad((T_new * Wt).sum(), y, retain_graph=True)
        out[f"gy_{name}"] = gy.numpy().copy()
        # every loss, with the gradient flowing back to y through the pose update
        with contextlib.redirect_stdout(io.StringIO()):
            losses = {
                "l1": ls.compute_ADD_L1_loss(T_gt, T_new, pts),
                "l1_ps": ls.compute_ADD_L1_loss(T_gt, T_new, pts, use_batch_mean=False),
                "l2": ls.compute_ADD_L2_loss(T_gt, T_new, pts),
                "l2_ps": ls.compute_ADD_L2_loss(T_gt, T_new, pts, use_batch_mean=False),
                "dis": ls.compute_disentangled_ADD_L1_loss(T_new, T_gt, pts),
                "dis_swapped": ls.compute_disentangled_ADD_L1_loss(T_gt, T_new, pts),   # train_model.py:280 (Q7)
                "scaled": ls.compute_scaled_disentl_ADD_L1_loss(T_prev, T_new, T_gt, pts),
            }
        for k, v in losses.items():
            out[f"{k}_{name}"] = v.detach().numpy().copy()
            if v.dim() == 0:
                (gl,) = torch.autograd.grad(v, y, retain_graph=True)
                out[f"g_{k}_{name}"] = gl.numpy().copy()
    # standalone rotation heads + transform_pts
    v3 = torch.tensor([[np.pi / 2, 0, 0], [0, 0, 0], [0.003, -0.004, 0.001], [1.0, -2.0, 0.5]], dtype=torch.float32)
    out["v3"] = v3.numpy()
    out["R_v3"] = rr.vec_3d_to_SO3(v3).numpy()        # rotation_representation.py:187-194 smoke vectors
    Tm = torch.stack([random_pose(g, 3) for _ in range(B)])     # (B,3,4,4)
    out["T_multi"] = Tm.numpy()
    out["xf_multi"] = ls.transform_pts(Tm, pts).numpy()
    out["xf_single"] = ls.transform_pts(T_gt, pts).numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, len(out), "arrays")


if __name__ == "__main__":
    main()
