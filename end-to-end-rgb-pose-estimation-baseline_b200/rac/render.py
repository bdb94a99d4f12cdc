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


_ids_cache = {}


def _resolve_meshes(meshes, B):
    """-> (ctypes int array of B handles, max triangle count).  One cache lookup per DISTINCT mesh of the batch (the
    reference passes B path strings, usually all equal: data_loaders.py:100-118)."""
    if not isinstance(meshes, (list, tuple)):
        meshes = [meshes] * B
    assert len(meshes) == B
    seen = {}
    handles = []
    for m in meshes:
        k = m if isinstance(m, str) else id(m)
        dm = seen.get(k)
        if dm is None:
            dm = seen[k] = device_mesh(m)
        handles.append(dm.handle)
    maxT = max(dm.T for dm in seen.values())
    if len(seen) == 1:
        key = (handles[0], B)
        ids = _ids_cache.get(key)
        if ids is None:
            if len(_ids_cache) > 256:
                _ids_cache.clear()
            ids = _ids_cache[key] = (ctypes.c_int * B)(*handles)
        return ids, maxT
    return (ctypes.c_int * B)(*handles), maxT


_k_cache = []          # [(weakref to the host tensor, data_ptr, version, device index, device copy)]


def _intrinsics_on_device(K, dev):
    """The drivers pass the same CPU `cam_mats` tensor in every call (train_model.py:240, Q5): its device copy is kept,
    keyed by the tensor object, its storage pointer and its in-place version counter."""
    if isinstance(K, torch.Tensor) and not K.is_cuda:
        for e in _k_cache:
            if e[0]() is K and e[1] == K.data_ptr() and e[2] == K._version and e[3] == dev.index:
                return e[4]
        Kt = K.to(device=dev, dtype=torch.float32).contiguous()
        import weakref
        _k_cache[:] = [e for e in _k_cache if e[0]() is not None and e[0]() is not K][-3:]
        _k_cache.append((weakref.ref(K), K.data_ptr(), K._version, dev.index, Kt))
        return Kt
    return torch.as_tensor(K).to(device=dev, dtype=torch.float32).contiguous()


def _prep(T_CO, K, meshes):
    dev = _lib.require_cuda()
    T = torch.as_tensor(T_CO).to(device=dev, dtype=torch.float32).contiguous()
    Kt = _intrinsics_on_device(K, dev)
    B = T.shape[0]
    assert T.shape == (B, 4, 4)
    assert Kt.shape == (B, 3, 3)
    ids, maxT = _resolve_meshes(meshes, B)
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


# ---- host gt crops: chunked upload on a copy stream, overlapped with the render ------------------------------
# The reference's caller hands prepare_model_input HOST arrays (train_model.py:243).  The gt crop is the one large
# input of a step (12 B/px); a plain `.to(device)` would queue its H2D copy behind the rasteriser on the current
# stream and block the host.  Instead the copy runs in chunks of images on a library-owned copy stream while the
# current stream is still rasterising, and each chunk is normalised + concatenated as soon as it has landed.
_copy_streams = {}
_gt_rings = {}
_GT_RING = 2
_GT_CHUNKS = 4


def _copy_stream(dev):
    s = _copy_streams.get(dev.index)
    if s is None:
        s = torch.cuda.Stream(device=dev)
        _copy_streams[dev.index] = s
    return s


def _gt_staging(dev, shape, dtype=torch.float32):
    """Ring of device staging buffers for uploaded gt crops; each entry carries the event of its last consumer."""
    key = (dev.index, tuple(shape), dtype)
    ring = _gt_rings.get(key)
    if ring is None:
        ring = {"next": 0, "bufs": [None] * _GT_RING, "done": [None] * _GT_RING}
        _gt_rings[key] = ring
    i = ring["next"]
    ring["next"] = (i + 1) % _GT_RING
    if ring["bufs"][i] is None:
        ring["bufs"][i] = torch.empty(shape, dtype=dtype, device=dev)
    return ring, i


def _overlappable(t, dtype=torch.float32):
    return (isinstance(t, torch.Tensor) and t.device.type == "cpu" and t.dtype == dtype and t.is_contiguous()
            and t.dim() == 4 and t.is_pinned() and not torch.cuda.is_current_stream_capturing())


def _assemble_call(L, init, gt, d, B, W, H, use_norm_depth, x, dev):
    need = L.rac_assemble_workspace_bytes(B)
    ws = _workspace(dev, need)
    base = (ws.data_ptr() + 255) // 256 * 256
    rc = L.rac_assemble_input(_lib.ptr(init), _lib.ptr(gt), _lib.ptr(d), B, W, H, int(bool(use_norm_depth)),
                              _lib.ptr(x), ctypes.c_void_p(base), ws.numel() - (base - ws.data_ptr()),
                              _lib.stream_ptr())
    _lib.check(rc)


# ---- normalised gt planes kept across the refinement iterations of a batch ----------------------------------
# The reference normalises and re-uploads the SAME gt crop in every refinement iteration of a batch
# (train_model.py:234-250: `gt_imgs` is drawn once, prepare_model_input runs inside the `for j` loop).  For torch
# tensors the (object, storage pointer, in-place version counter) triple identifies unchanged content safely, so the
# normalised planes of the last few crops stay on the device and later iterations skip the upload and the
# normalisation.  numpy arrays carry no version counter and are never cached.
_gt_cache = []          # [(weakref, data_ptr, version, shape, device index, planes (B,3,H,W))], most recent last
_GT_CACHE_ENTRIES = 2
_gt_cache_enabled = True


def set_gt_cache(enabled):
    """Switch the cross-iteration gt cache on/off (bench.py's per-step end-to-end figure switches it off: every
    step must pay its own upload); returns the previous setting."""
    global _gt_cache_enabled
    prev = _gt_cache_enabled
    _gt_cache_enabled = bool(enabled)
    if not enabled:
        _gt_cache.clear()
    return prev


def _gt_cache_lookup(t, dev):
    if not (_gt_cache_enabled and isinstance(t, torch.Tensor)):
        return None
    for e in _gt_cache:
        if e[0]() is t and e[1] == t.data_ptr() and e[2] == t._version and e[3] == tuple(t.shape) and e[4] == dev.index:
            return e[5]
    return None


def _gt_cache_store(t, dev, planes):
    if not (_gt_cache_enabled and isinstance(t, torch.Tensor)):
        return
    import weakref
    _gt_cache[:] = [e for e in _gt_cache if e[0]() is not None and e[0]() is not t][-(_GT_CACHE_ENTRIES - 1):]
    _gt_cache.append((weakref.ref(t), t.data_ptr(), t._version, tuple(t.shape), dev.index, planes.clone()))


# ---- 8-bit gt crops ---------------------------------------------------------------------------------------------
# A uint8 (B,H,W,3) crop is taken as the raw `real.png` codes: the `/ 255.0` of image_dataloaders.py:79-83 moves into
# rac_gt_norm_u8, and the crop crosses PCIe at a quarter of the bytes of the f32 form.
def _is_u8(t):
    return (isinstance(t, torch.Tensor) and t.dtype == torch.uint8) or (isinstance(t, np.ndarray) and t.dtype == np.uint8)


def gt_norm_u8_device(gt_u8, x, use_norm_depth=False):
    """normalize_img(gt_u8 / 255.0) -> the three gt planes of x (B,C,H,W) f32 CUDA; gt_u8 (B,H,W,3) uint8, host or device.
    A pinned host tensor is uploaded on the library's copy stream (it overlaps whatever the current stream is running)."""
    L = _lib.lib()
    dev = x.device
    B, C, H, W = x.shape
    assert C == (7 if use_norm_depth else 6) and tuple(gt_u8.shape) == (B, H, W, 3) and x.is_contiguous() and x.dtype == torch.float32
    ring = None
    if _overlappable(gt_u8, torch.uint8):
        main = torch.cuda.current_stream(dev)
        cs = _copy_stream(dev)
        ring, slot = _gt_staging(dev, (B, H, W, 3), torch.uint8)
        g = ring["bufs"][slot]
        if ring["done"][slot] is not None:
            cs.wait_event(ring["done"][slot])
        with torch.cuda.stream(cs):
            g.copy_(gt_u8, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
        main.wait_event(ev)
    else:
        g = torch.as_tensor(gt_u8).to(device=dev).contiguous()
    need = L.rac_gt_u8_workspace_bytes(B)
    ws = _workspace(dev, need + 256)
    base = (ws.data_ptr() + 255) // 256 * 256
    rc = L.rac_gt_norm_u8(_lib.ptr(g), B, W, H, int(bool(use_norm_depth)), _lib.ptr(x), ctypes.c_void_p(base),
                          ws.numel() - (base - ws.data_ptr()), _lib.stream_ptr())
    _lib.check(rc)
    if ring is not None:
        done = torch.cuda.Event()
        done.record(torch.cuda.current_stream(dev))
        ring["done"][slot] = done
    return x


def assemble_input_device(init_hwc, gt_hwc, depth, use_norm_depth=False):
    """prepare_model_input on renders that already exist (iteration 0): -> x (B,C,H,W) f32 CUDA."""
    L = _lib.lib()
    dev = _lib.require_cuda()
    init = torch.as_tensor(init_hwc).to(device=dev, dtype=torch.float32).contiguous()
    B, H, W, _ = init.shape
    assert init.shape == (B, H, W, 3) and tuple(gt_hwc.shape) == (B, H, W, 3)
    d = None
    if use_norm_depth:
        d = torch.as_tensor(depth).to(device=dev, dtype=torch.float32).contiguous()
        assert d.shape == (B, H, W)
    C = 7 if use_norm_depth else 6
    x = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    cached = _gt_cache_lookup(gt_hwc, dev)
    if cached is not None:
        x[:, C - 3:].copy_(cached)
        _assemble_call(L, init, None, d, B, W, H, use_norm_depth, x, dev)
        return x
    if _is_u8(gt_hwc):
        gt_norm_u8_device(gt_hwc, x, use_norm_depth)
        _assemble_call(L, init, None, d, B, W, H, use_norm_depth, x, dev)
        _gt_cache_store(gt_hwc, dev, x[:, C - 3:])
        return x
    if _overlappable(gt_hwc) and B >= 2 * _GT_CHUNKS:
        main = torch.cuda.current_stream(dev)
        cs = _copy_stream(dev)
        ring, slot = _gt_staging(dev, (B, H, W, 3))
        gt = ring["bufs"][slot]
        if ring["done"][slot] is not None:
            cs.wait_event(ring["done"][slot])          # the staging buffer's previous consumer has finished
        per = (B + _GT_CHUNKS - 1) // _GT_CHUNKS
        events = []
        with torch.cuda.stream(cs):
            for b0 in range(0, B, per):
                b1 = min(B, b0 + per)
                gt[b0:b1].copy_(gt_hwc[b0:b1], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                events.append((b0, b1, ev))
        for b0, b1, ev in events:
            main.wait_event(ev)
            _assemble_call(L, init[b0:b1], gt[b0:b1], d[b0:b1] if d is not None else None, b1 - b0, W, H,
                           use_norm_depth, x[b0:b1], dev)
        done = torch.cuda.Event()
        done.record(main)
        ring["done"][slot] = done
        _gt_cache_store(gt_hwc, dev, x[:, C - 3:])
        return x
    gt = torch.as_tensor(gt_hwc).to(device=dev, dtype=torch.float32).contiguous()
    _assemble_call(L, init, gt, d, B, W, H, use_norm_depth, x, dev)
    _gt_cache_store(gt_hwc, dev, x[:, C - 3:])
    return x


def to_numpy(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
