"""Drop-in for the reference's rotation_representation.py.

Same public names and signatures (ref: rotation_representation.py:4-19, 22-39, 41-72, 74-106,
108-118, 131-184); each function is one fused CUDA kernel (csrc/pose_update.cu) with a
hand-written backward, instead of ~25 small torch kernels plus a batched cuSOLVER SVD.
"""
import torch

from rac import _lib
from rac.ops import PoseUpdateFn, RotationFn, rot_id


def compute_rotation_matrix_from_ortho6d(poses):
    """ref: rotation_representation.py:4-19 — Gram-Schmidt, columns (x, y, z), no epsilon (Q9)."""
    assert poses.shape[-1] == 6
    lead = poses.shape[:-1]
    R = RotationFn.apply(poses.reshape(-1, 6), _lib.RAC_ROT["6D"], 0.0)
    return R.reshape(*lead, 3, 3)


def symmetric_orthogonalization(x):
    """ref: rotation_representation.py:22-39 — SVD projection of a (B,9) vector onto SO(3)."""
    return RotationFn.apply(x.reshape(-1, 9), _lib.RAC_ROT["SVD"], 0.0)


def so3_exp_map(log_rot, eps: float = 0.0001):
    """ref: rotation_representation.py:41-72 — Rodrigues with theta^2 clamped at eps (Q8)."""
    _, dim = log_rot.shape
    if dim != 3:
        raise ValueError("Input tensor shape has to be Nx3.")
    return RotationFn.apply(log_rot, _lib.RAC_ROT["3D"], eps)


def hat(v):
    """ref: rotation_representation.py:74-106 — (N,3) -> (N,3,3) skew-symmetric matrices."""
    N, dim = v.shape
    if dim != 3:
        raise ValueError("Input vectors have to be 3-dimensional.")
    x, y, z = v.unbind(1)
    o = torch.zeros_like(x)
    return torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=1).reshape(N, 3, 3)


def vec_3d_to_SO3(x):
    """ref: rotation_representation.py:108-118."""
    bsz = x.shape[0]
    assert x.shape == (bsz, 3)
    return so3_exp_map(x)


def calculate_T_CO_pred(model_output, T_CO_init, rot_repr, Ks):
    """ref: rotation_representation.py:131-184.  Keeps Q1 (y reads the x translation), Q2 (first
    sample's intrinsics for the whole batch), Q3 (f32 output on model_output.device), Q6 (f64 pose
    input keeps the translation update in f64)."""
    bsz = model_output.shape[0]
    if rot_repr == 'SVD':
        assert model_output.shape == (bsz, 12)
    elif rot_repr == '6D':
        assert model_output.shape == (bsz, 9)
    elif rot_repr == '3D':
        assert model_output.shape == (bsz, 6)
    assert T_CO_init.shape == (bsz, 4, 4)
    assert Ks.shape == (bsz, 3, 3)
    rid = rot_id(rot_repr)
    K0 = Ks[0]
    fx, fy = float(K0[0, 0]), float(K0[1, 1])
    if not isinstance(T_CO_init, torch.Tensor):
        T_CO_init = torch.as_tensor(T_CO_init)
    return PoseUpdateFn.apply(model_output, T_CO_init, rid, fx, fy)
