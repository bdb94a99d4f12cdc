"""Pins oracle/pose_loss_ref.py (steps 3-5) against (a) golden vectors produced by executing the
reference and (b) the reference itself when /root/reference is present (build container only)."""
import contextlib
import io
import os
import sys

import numpy as np
import pytest
import torch

from oracle import pose_loss_ref as R

GOLD = os.path.join(os.path.dirname(__file__), "golden", "pose_loss_golden.npz")
HAVE_REF = os.path.isdir("/root/reference")


@pytest.fixture(scope="module")
def gold():
    return {k: v for k, v in np.load(GOLD).items()}


def t(a):
    return torch.from_numpy(np.array(a))


@pytest.mark.parametrize("rot", ["6D", "SVD", "3D"])
def test_pose_update_matches_golden(gold, rot):
    y, T_in, Ks = t(gold[f"y_{rot}"]), t(gold["T_in"]), t(gold["Ks"])
    out = R.pose_update(y, T_in, rot, Ks)
    np.testing.assert_allclose(out.numpy(), gold[f"Tnew_{rot}_f32"], rtol=1e-5, atol=1e-6)
    out64 = R.pose_update(y, T_in.double(), rot, Ks)
    np.testing.assert_allclose(out64.numpy(), gold[f"Tnew_{rot}_f64"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("rot", ["6D", "SVD", "3D"])
def test_losses_and_grads_match_golden(gold, rot):
    y = t(gold[f"y_{rot}"]).requires_grad_(True)
    T_in, T_gt, T_prev, pts, Ks = (t(gold[k]) for k in ("T_in", "T_gt", "T_prev", "pts", "Ks"))
    T_new = R.pose_update(y, T_in, rot, Ks)
    (gy,) = torch.autograd.grad((T_new * t(gold[f"W_{rot}"])).sum(), y, retain_graph=True)
    np.testing.assert_allclose(gy.numpy(), gold[f"gy_{rot}"], rtol=2e-4, atol=1e-5)
    losses = {
        "l1": R.add_l1(T_gt, T_new, pts), "l1_ps": R.add_l1(T_gt, T_new, pts, False),
        "l2": R.add_l2(T_gt, T_new, pts), "l2_ps": R.add_l2(T_gt, T_new, pts, False),
        "dis": R.disentangled(T_new, T_gt, pts), "dis_swapped": R.disentangled(T_gt, T_new, pts),
        "scaled": R.disentangled_scaled(T_prev, T_new, T_gt, pts),
    }
    for k, v in losses.items():
        np.testing.assert_allclose(v.detach().numpy(), gold[f"{k}_{rot}"], rtol=1e-5, atol=1e-7, err_msg=k)
        if v.dim() == 0:
            (g,) = torch.autograd.grad(v, y, retain_graph=True)
            np.testing.assert_allclose(g.numpy(), gold[f"g_{k}_{rot}"], rtol=2e-4, atol=1e-6, err_msg=k)


def test_expmap_known_answers(gold):
    """rotation_representation.py:187-194: Rx(90 deg) and the clamped zero vector (exactly I)."""
    Rm = R.rot_exp(t(gold["v3"]))
    np.testing.assert_allclose(Rm.numpy(), gold["R_v3"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(Rm[0].numpy(), [[1, 0, 0], [0, 0, -1], [0, 1, 0]], atol=1e-6)
    np.testing.assert_array_equal(Rm[1].numpy(), np.eye(3, dtype=np.float32))


def test_transform_pts_matches_golden(gold):
    np.testing.assert_allclose(R.xform(t(gold["T_multi"]), t(gold["pts"])).numpy(), gold["xf_multi"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(R.xform(t(gold["T_gt"]), t(gold["pts"])).numpy(), gold["xf_single"], rtol=1e-5, atol=1e-6)


@pytest.mark.skipif(not HAVE_REF, reason="/root/reference only exists in the build container")
@pytest.mark.parametrize("rot,D", [("6D", 9), ("SVD", 12), ("3D", 6)])
def test_restatement_equals_reference_on_fresh_inputs(rot, D):
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    from make_pose_loss_golden import load_reference, random_pose
    rr, ls = load_reference()
    g = torch.Generator().manual_seed(99)
    B, N = 16, 200
    y = 0.3 * torch.randn(B, D, generator=g)
    y[:, D - 1] += 1.0
    y.requires_grad_(True)
    T_in, T_gt, T_prev = random_pose(g, B), random_pose(g, B), random_pose(g, B)
    pts = torch.randn(B, N, 3, generator=g) * 0.3
    Ks = torch.tensor([[444.44, 0, 160.0], [0, 444.44, 160.0], [0, 0, 1]]).repeat(B, 1, 1)
    a = rr.calculate_T_CO_pred(y, T_in, rot, Ks)
    b = R.pose_update(y, T_in, rot, Ks)
    np.testing.assert_allclose(b.detach().numpy(), a.detach().numpy(), rtol=1e-5, atol=1e-6)
    with contextlib.redirect_stdout(io.StringIO()):
        pairs = [(ls.compute_ADD_L1_loss(T_gt, a, pts), R.add_l1(T_gt, b, pts)),
                 (ls.compute_ADD_L2_loss(T_gt, a, pts), R.add_l2(T_gt, b, pts)),
                 (ls.compute_disentangled_ADD_L1_loss(a, T_gt, pts), R.disentangled(b, T_gt, pts)),
                 (ls.compute_scaled_disentl_ADD_L1_loss(T_prev, a, T_gt, pts), R.disentangled_scaled(T_prev, b, T_gt, pts))]
    for ra, rb in pairs:
        np.testing.assert_allclose(rb.item(), ra.item(), rtol=1e-5)
        (ga,) = torch.autograd.grad(ra, y, retain_graph=True)
        (gb,) = torch.autograd.grad(rb, y, retain_graph=True)
        np.testing.assert_allclose(gb.numpy(), ga.numpy(), rtol=5e-4, atol=1e-6)
