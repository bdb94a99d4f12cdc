"""Device-side entry points for steps 1-2: batched render and render fused with input assembly.

These are the calls the shadow modules (renderer.py, data_loaders.py) route through; they take
and return CUDA tensors so a caller that keeps its loop on the device pays no host transfer.
ref: data_loaders.py:84-118 (render_batch), data_loaders.py:158-177 (prepare_model_input).
"""
import ctypes

import numpy as np
import torch

from . import _lib
from .mesh import device_mesh

_ws_cache = {}


def _workspace(dev, nbytes):
    """Grow-only per-device scratch.  Reuse is stream-ordered: every kernel of a call is enqueued on
    the current stream, so a later call on the same stream cannot overtake an earlier one."""
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=dev)
        _ws_cache[key] = ws
    return ws


def _scene_struct(scene):
    if scene is None:
        return None
    s = _lib.RacScene()
    _lib.lib().rac_scene_defaults(ctypes.byref(s))
    for k, v in scene.items():
        setattr(s, k, float(v))
    return ctypes.byref(s)


def _prep(T_CO, K, meshes):
    dev = _lib.require_cuda()
    T = torch.as_tensor(T_CO).to(device=dev, dtype=torch.float32).contiguous()
    Kt = torch.as_tensor(K).to(device=dev, dtype=torch.float32).contiguous()
    B = T.shape[0]
    assert T.shape == (B, 4, 4)
    assert Kt.shape == (B, 3, 3)
    if not isinstance(meshes, (list, tuple)):
        meshes = [meshes] * B
    assert len(meshes) == B
    dms = [device_mesh(m) for m in meshes]
    ids = (ctypes.c_int * B)(*[d.handle for d in dms])
    maxT = max(d.T for d in dms)
    return dev, T, Kt, B, ids, maxT


def _wh(img_size):
    """img_size: int (square, the reference's only case, renderer.py:78) or (W, H)."""
    if isinstance(img_size, (tuple, list)):
        return int(img_size[0]), int(img_size[1])
    return int(img_size), int(img_size)


def render_batch_device(T_CO, meshes, K, img_size, scene=None, want_rgb=True, want_depth=True, want_rgba8=False):
    """-> dict(rgb (B,H,W,3) f32, depth (B,H,W) f32, rgba8 (B,H,W,4) u8) of CUDA tensors."""
    L = _lib.lib()
    dev, T, Kt, B, ids, maxT = _prep(T_CO, K, meshes)
    W, H = _wh(img_size)
    need = L.rac_render_workspace_bytes(B, W, H, maxT)
    ws = _workspace(dev, need)
    out = {}
    rgb = torch.empty((B, H, W, 3), dtype=torch.float32, device=dev) if want_rgb else None
    depth = torch.empty((B, H, W), dtype=torch.float32, device=dev) if want_depth else None
    rgba8 = torch.empty((B, H, W, 4), dtype=torch.uint8, device=dev) if want_rgba8 else None
    base = (ws.data_ptr() + 255) // 256 * 256
    rc = L.rac_render_batch(ids, _lib.ptr(T), _lib.ptr(Kt), B, W, H, _scene_struct(scene), _lib.ptr(rgb),
                            _lib.ptr(depth), _lib.ptr(rgba8), ctypes.c_void_p(base),
                            ws.numel() - (base - ws.data_ptr()), _lib.stream_ptr())
    _lib.check(rc)
    out["rgb"], out["depth"], out["rgba8"] = rgb, depth, rgba8
    return out


def render_concat_device(T_CO, meshes, K, img_size, gt_hwc, use_norm_depth=False, scene=None,
                         want_depth=False, want_rgb=False, out=None):
    """Fused render + prepare_model_input: -> x (B,C,H,W) f32 CUDA tensor [, depth, rgb]."""
    L = _lib.lib()
    dev, T, Kt, B, ids, maxT = _prep(T_CO, K, meshes)
    W, H = _wh(img_size)
    gt = torch.as_tensor(gt_hwc).to(device=dev, dtype=torch.float32).contiguous()
    assert gt.shape == (B, H, W, 3)
    C = 7 if use_norm_depth else 6
    need = L.rac_render_workspace_bytes(B, W, H, maxT)
    ws = _workspace(dev, need)
    x = out if out is not None else torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    assert x.shape == (B, C, H, W) and x.is_contiguous() and x.dtype == torch.float32
    depth = torch.empty((B, H, W), dtype=torch.float32, device=dev) if want_depth else None
    rgb = torch.empty((B, H, W, 3), dtype=torch.float32, device=dev) if want_rgb else None
    base = (ws.data_ptr() + 255) // 256 * 256
    rc = L.rac_render_concat(ids, _lib.ptr(T), _lib.ptr(Kt), B, W, H, _scene_struct(scene), _lib.ptr(gt),
                             int(bool(use_norm_depth)), _lib.ptr(x), _lib.ptr(depth), _lib.ptr(rgb),
                             ctypes.c_void_p(base), ws.numel() - (base - ws.data_ptr()), _lib.stream_ptr())
    _lib.check(rc)
    return x, depth, rgb


def assemble_input_device(init_hwc, gt_hwc, depth, use_norm_depth=False):
    """prepare_model_input on renders that already exist (iteration 0): -> x (B,C,H,W) f32 CUDA."""
    L = _lib.lib()
    dev = _lib.require_cuda()
    init = torch.as_tensor(init_hwc).to(device=dev, dtype=torch.float32).contiguous()
    gt = torch.as_tensor(gt_hwc).to(device=dev, dtype=torch.float32).contiguous()
    B, H, W, _ = init.shape
    assert init.shape == (B, H, W, 3) and gt.shape == (B, H, W, 3)
    d = None
    if use_norm_depth:
        d = torch.as_tensor(depth).to(device=dev, dtype=torch.float32).contiguous()
        assert d.shape == (B, H, W)
    C = 7 if use_norm_depth else 6
    x = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    need = L.rac_assemble_workspace_bytes(B)
    ws = _workspace(dev, need)
    base = (ws.data_ptr() + 255) // 256 * 256
    rc = L.rac_assemble_input(_lib.ptr(init), _lib.ptr(gt), _lib.ptr(d), B, W, H, int(bool(use_norm_depth)),
                              _lib.ptr(x), ctypes.c_void_p(base), ws.numel() - (base - ws.data_ptr()),
                              _lib.stream_ptr())
    _lib.check(rc)
    return x


def to_numpy(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
