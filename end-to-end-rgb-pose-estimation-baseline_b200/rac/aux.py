"""Device-side entry points for the two rows next to the hot path:
N3 batched pose sampling (ref: se3_helpers.py:76-88, data_loaders.py:58-67) and
N4 silhouette ring / overlay blend (ref: visualization.py:19-41).  CUDA only, like the rest."""
import ctypes

import numpy as np
import torch

from . import _lib


def sample_poses_device(batch_size, scene_config, generator=None, variates=None):
    """-> (T_CO_init, T_CO_gt), two (B,4,4) f32 CUDA tensors drawn like `get_T_CO_init_and_gt` (se3_helpers.py:76-88),
    one kernel launch for the batch instead of a Python loop (data_loaders.py:58-67).
    `variates` = (u01 (B,12), n01 (B,3)) overrides the RNG (tests feed the numpy restatement's draws)."""
    L = _lib.lib()
    dev = _lib.require_cuda()
    B = int(batch_size)
    if variates is None:
        u = torch.rand((B, 12), device=dev, dtype=torch.float32, generator=generator)
        n = torch.randn((B, 3), device=dev, dtype=torch.float32, generator=generator)
    else:
        u = torch.as_tensor(variates[0]).to(device=dev, dtype=torch.float32).contiguous()
        n = torch.as_tensor(variates[1]).to(device=dev, dtype=torch.float32).contiguous()
    assert u.shape == (B, 12) and n.shape == (B, 3)
    T_init = torch.empty((B, 4, 4), dtype=torch.float32, device=dev)
    T_gt = torch.empty((B, 4, 4), dtype=torch.float32, device=dev)
    rc = L.rac_sample_poses(_lib.ptr(u), _lib.ptr(n), B,
                            float(scene_config["distance_cam_to_world"]),
                            float(scene_config["distance_cam_to_world_deviation"]),
                            float(scene_config["world_to_object_gt_transl_deviation"]),
                            float(scene_config["world_to_object_transl_deviation"]),
                            float(scene_config["world_to_object_angle_deviation"]),
                            _lib.ptr(T_init), _lib.ptr(T_gt), _lib.stream_ptr())
    _lib.check(rc)
    return T_init, T_gt


def _batched(img):
    t = img.device_tensor if hasattr(img, "device_tensor") else torch.as_tensor(img)
    single = t.dim() == 3
    if single:
        t = t.unsqueeze(0)
    assert t.dim() == 4 and t.shape[-1] == 3, "image must be (H,W,3) or (B,H,W,3)"
    return t, single


def silhouette_device(img, silhouette_only=False, colour=(1.0, 0.0, 0.0)):
    """create_silhouette for one image or a batch: -> ring mask (…,H,W) or painted image (…,H,W,3), f32 CUDA."""
    L = _lib.lib()
    dev = _lib.require_cuda()
    t, single = _batched(img)
    t = t.to(device=dev, dtype=torch.float32).contiguous()
    B, H, W, _ = t.shape
    out = torch.empty((B, H, W) if silhouette_only else (B, H, W, 3), dtype=torch.float32, device=dev)
    col = (ctypes.c_float * 3)(*[float(c) for c in colour])
    _lib.check(L.rac_silhouette(_lib.ptr(t), B, W, H, int(bool(silhouette_only)), col, _lib.ptr(out), _lib.stream_ptr()))
    return out[0] if single else out


def _codes_or_float(a, dev):
    """f64 host input: quantise on the host in f64 exactly as np.uint8(im * 255.0) does; anything else goes to the
    device as f32 and is quantised by the kernel in f32 (what numpy does for an f32 array)."""
    if isinstance(a, np.ndarray) and a.dtype == np.float64:
        return torch.as_tensor(np.uint8(a * 255.0).astype(np.float32)).to(dev), True
    t = a.device_tensor if hasattr(a, "device_tensor") else torch.as_tensor(a)
    if t.dtype == torch.float64:
        return torch.as_tensor(np.uint8(t.cpu().numpy() * 255.0).astype(np.float32)).to(dev), True
    return t.to(device=dev, dtype=torch.float32).contiguous(), False


def blend_device(im1, im2, alpha):
    """blend_imgs: -> f32 CUDA tensor of the inputs' shape."""
    L = _lib.lib()
    dev = _lib.require_cuda()
    a, qa = _codes_or_float(im1, dev)
    b, qb = _codes_or_float(im2, dev)
    assert a.shape == b.shape
    if qa != qb:      # mixed precisions: bring the f32 side to codes too (f32 arithmetic, as numpy would)
        if not qa:
            a = torch.floor(a * 255.0).clamp_(0, 255)
        else:
            b = torch.floor(b * 255.0).clamp_(0, 255)
        qa = qb = True
    out = torch.empty_like(a)
    _lib.check(L.rac_blend(_lib.ptr(a), _lib.ptr(b), float(alpha), int(qa), a.numel(), _lib.ptr(out), _lib.stream_ptr()))
    return out
