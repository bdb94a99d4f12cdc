"""Drop-in for the reference's loss.py.

Same public names and signatures (ref: loss.py:5-18, 20-33, 35-52, 60-78, 81-97, 101-122);
every loss is one fused transform-and-reduce CUDA kernel (csrc/add_loss.cu) whose forward also
produces d(loss)/dT, so autograd's backward is a 16-float scale per sample.
"""
import numpy as np
import torch

from rac import _lib
from rac.ops import AddLossFn, TransformPtsFn
from rotation_representation import calculate_T_CO_pred   # noqa: F401  (ref: loss.py:2)


def _check(TCO_gt, TCO_pred, points):
    bsz = len(TCO_gt)
    assert TCO_pred.shape == (bsz, 4, 4) and TCO_gt.shape == (bsz, 4, 4)
    assert points.dim() == 3 and points.shape[-1] == 3


def compute_ADD_L1_loss(TCO_gt, TCO_pred, points, use_batch_mean=True):
    """ref: loss.py:5-18."""
    _check(TCO_gt, TCO_pred, points)
    return AddLossFn.apply(TCO_gt, TCO_pred, None, points, _lib.LOSS_ADD_L1, bool(use_batch_mean))


def compute_ADD_L2_loss(TCO_gt, TCO_pred, points, use_batch_mean=True):
    """ref: loss.py:20-33 (the reference's two debug shape prints are kept)."""
    _check(TCO_gt, TCO_pred, points)
    print(torch.Size((points.shape[0], points.shape[1], 3)))
    out = AddLossFn.apply(TCO_gt, TCO_pred, None, points, _lib.LOSS_ADD_L2, bool(use_batch_mean))
    print(torch.Size((points.shape[0],)))
    return out


def compute_angular_dist(T_CO_gt, T_CO_pred):
    """ref: loss.py:35-43 — geodesic angle of sample 0 only; arccos is not clamped (NaN possible).
    The value is computed in the input precision exactly as the reference does and then handed back as numpy float64.
    test_model.py:196,208,277 multiplies it by 180/np.pi and json.dump()s it: with the float32 poses the loader produces
    (image_dataloaders.py:75) that product was a float64 under the NumPy 1.x scalar rules the reference was written
    for, and stays a float32 — which json refuses — under NumPy >= 2.  Widening here gives the unchanged driver the
    NumPy 1.x result on either version (tests/test_drivers_unchanged.py::test_test_model_unchanged)."""
    R_gt = T_CO_gt[0, :3, :3].detach().cpu().numpy()
    R_pred = T_CO_pred[0, :3, :3].detach().cpu().numpy()
    return np.float64(np.arccos((np.trace(R_pred @ R_gt.T) - 1.0) / 2.0))


def compute_transl_dist(T_CO_gt, T_CO_pred):
    """ref: loss.py:45-52 — translation distance of sample 0 only."""
    d = T_CO_gt[0, :3, 3].detach().cpu().numpy() - T_CO_pred[0, :3, 3].detach().cpu().numpy()
    return np.sqrt(np.sum(np.square(d)))


def transform_pts(T, pts):
    """ref: loss.py:60-78.  T (B,4,4) or (B,M,4,4); pts (B,N,3)."""
    bsz = T.shape[0]
    n_pts = pts.shape[1]
    assert pts.shape == (bsz, n_pts, 3)
    if T.dim() == 4:
        assert T.shape[-2:] == (4, 4)
    elif T.dim() == 3:
        assert T.shape == (bsz, 4, 4)
    else:
        raise ValueError('Unsupported shape for T', T.shape)
    return TransformPtsFn.apply(T, pts)


def compute_disentangled_ADD_L1_loss(T_CO_pred, T_CO_gt, points):
    """ref: loss.py:81-97.  Note train_model.py:280 calls this with (gt, pred) swapped (Q7); both
    arguments are differentiable here, as they are in the reference."""
    _check(T_CO_gt, T_CO_pred, points)
    return AddLossFn.apply(T_CO_pred, T_CO_gt, None, points, _lib.LOSS_DISENTANGLED, True)


def compute_scaled_disentl_ADD_L1_loss(T_CO_pred_prev, T_CO_pred, T_CO_gt, points):
    """ref: loss.py:101-122."""
    _check(T_CO_gt, T_CO_pred, points)
    return AddLossFn.apply(T_CO_pred, T_CO_gt, T_CO_pred_prev.detach(), points, _lib.LOSS_DISENT_SCALED, True)
