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
    """Grow-only scratch per (device, stream), at most 8 of them.  Reuse is stream-ordered: every kernel of a call is enqueued on
    the current stream, so a later call on the same stream cannot overtake an earlier one."""
    key = (dev.index, _lib.raw_stream(dev.index))
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        if ws is None:
            while len(_ws_cache) >= 8:      # short-lived streams (one per captured graph, say) must not pin a workspace each forever
                _ws_cache.pop(next(iter(_ws_cache)))
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
        while len(_gt_rings) >= 4:          # bounded like the pinned rings (the caching allocator reuses the freed blocks)
            _gt_rings.pop(next(iter(_gt_rings)))
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
# (train_model.py:234-250: `gt_imgs` is drawn once, prepare_model_input runs inside the `for j` loop).  The normalised
# planes of the last few crops stay on the device and later iterations skip the upload and the normalisation
# (SURVEY.md §8f N1).  The reference itself recomputes on every call (data_loaders.py:163-166), so a hit must mean
# "the same object, holding what it held when the planes were made".  What counts as that:
#   host arrays (numpy — what the drivers pass, train_model.py:226 — and CPU tensors):
#     (1) IDENTITY: a weak reference to the very object that was passed.  A new array never hits, whatever address the
#         allocator gave it (the unchanged driver allocates a new `gt_imgs` per batch and freed addresses come back);
#     (2) CONTENT, against in-place refills of that object (through a numpy view, from_numpy-shared storage, ...; torch's
#         version counter does not see those): a fingerprint taken at insert and at every lookup — the raw bytes of one
#         pixel (all three channels) out of every _FP_LATTICE x _FP_LATTICE block of EVERY image, on a lattice whose
#         column offset shifts from row band to row band (no aliasing with the row pitch), plus the first and last 256
#         values.  ~0.4 ms at 64 x 320 x 240 (19 200 cache lines of pageable memory).
#     Guarantee: an in-place edit that changes every pixel of some 16 x 16 block of some image (a refilled buffer, a
#     replaced image, a pasted / augmented region of that size) is always a miss.  NOT guaranteed: edits smaller than
#     the lattice (single pixels, thin lines) between two calls on the same object — a caller that edits crops in place
#     at that granularity between refinement iterations switches the cache off: rac.render.set_gt_cache(False).
#   CUDA tensors: object identity, storage pointer and in-place version counter (no host copy of a fingerprint
#       without a sync; device buffers cannot be written behind torch's back through numpy).
_gt_cache = []          # [dict(key, val, ref, planes (B,3,H,W))], most recent last
_GT_CACHE_ENTRIES = 2
_gt_cache_enabled = True
_FP_LATTICE = 16
_FP_EDGE = 256
_fp_index = {}          # (B, H, W) -> flat pixel indices of the lattice (int64)


def set_gt_cache(enabled):
    """Switch the cross-iteration gt cache on/off; returns the previous setting.  Off = every prepare_model_input
    uploads and normalises its crop, exactly like the reference (use it when crops are edited in place at a finer
    granularity than the fingerprint lattice; bench.py's fresh-crop-every-step figure also switches it off)."""
    global _gt_cache_enabled
    prev = _gt_cache_enabled
    _gt_cache_enabled = bool(enabled)
    if not enabled:
        _gt_cache.clear()
    return prev


def _host_view(t):
    """numpy view (no copy) of a host array the cache / staging code can fingerprint, or None."""
    if isinstance(t, np.ndarray):
        return t if t.flags.c_contiguous else None
    if isinstance(t, torch.Tensor) and t.device.type == "cpu" and t.is_contiguous():
        return t.detach().numpy()
    return None


def _lattice(B, H, W):
    idx = _fp_index.get((B, H, W))
    if idx is None:
        s = _FP_LATTICE
        per = []
        for j, y in enumerate(range(min(s // 2, H - 1), H, s)):
            xs = np.arange((5 * j) % s, W, s, dtype=np.int64)      # 5 is coprime to 16: the offset walks all residues
            per.append(y * W + xs)
        per = np.concatenate(per)
        idx = (np.arange(B, dtype=np.int64)[:, None] * (H * W) + per[None]).reshape(-1)
        if len(_fp_index) >= 8:
            _fp_index.clear()
        _fp_index[(B, H, W)] = idx
    return idx


def _fingerprint(a):
    """a: C-contiguous (B,H,W,3) host array."""
    B, H, W, C = a.shape
    flat = a.reshape(-1)
    return np.take(a.reshape(B * H * W, C), _lattice(B, H, W), axis=0).tobytes() + flat[:_FP_EDGE].tobytes() + flat[-_FP_EDGE:].tobytes()


def _gt_cache_key(t, dev):
    """-> (key, validator) or None when `t` cannot be cached."""
    a = _host_view(t)
    if a is not None:
        if a.ndim != 4 or a.size == 0:
            return None
        return ("host", a.__array_interface__["data"][0], a.shape, a.dtype.str, a.strides, dev.index), _fingerprint(a)
    if isinstance(t, torch.Tensor) and t.is_cuda:
        return ("cuda", id(t), t.data_ptr(), tuple(t.shape), str(t.dtype), dev.index), t._version
    return None


def _gt_cache_lookup(t, dev, kv=None):
    """kv: a (key, validator) pair already computed by _gt_cache_key for this very call (the fingerprint of a host array
    costs a few tenths of a millisecond: take it once per prepare_model_input, not once per lookup and once per store)."""
    if not _gt_cache_enabled:
        return None
    kv = kv or _gt_cache_key(t, dev)
    if kv is None:
        return None
    for e in _gt_cache:
        if e["ref"]() is t and e["key"] == kv[0] and e["val"] == kv[1]:
            return e["planes"]
    return None


def _gt_cache_store(t, dev, planes, kv=None):
    if not _gt_cache_enabled:
        return
    kv = kv or _gt_cache_key(t, dev)
    if kv is None:
        return
    import weakref
    try:
        ref = weakref.ref(t)        # identity: an entry whose object has died can never hit (addresses and id()s are recycled)
    except TypeError:
        return
    keep = [e for e in _gt_cache if e["ref"]() is not None and e["ref"]() is not t]
    _gt_cache[:] = keep[-(_GT_CACHE_ENTRIES - 1):] if _GT_CACHE_ENTRIES > 1 else []
    _gt_cache.append({"key": kv[0], "val": kv[1], "ref": ref, "planes": planes.clone()})


# ---- pinned host staging (library-owned) ----------------------------------------------------------------------
# Pageable host arrays (what the unchanged loader produces) are first brought into a pinned ring buffer — converted on
# the way (f64 -> u8 codes by rac_host_pack_gt_u8, or f64 -> f32 by a threaded torch copy) — so that the H2D copy is
# asynchronous and runs at full PCIe rate.  A ring slot is reused only after the copy that read it has completed.
_host_rings = {}
_HOST_RING = 2
_MAX_STAGING_SHAPES = 4


def _host_staging(shape, dtype):
    key = (tuple(shape), dtype)
    ring = _host_rings.get(key)
    if ring is None:
        while len(_host_rings) >= _MAX_STAGING_SHAPES:      # ragged last batches bring new shapes: keep pinned memory bounded
            old = _host_rings.pop(next(iter(_host_rings)))
            for ev in old["ev"]:
                if ev is not None:
                    ev.synchronize()
        ring = {"next": 0, "bufs": [None] * _HOST_RING, "ev": [None] * _HOST_RING}
        _host_rings[key] = ring
    i = ring["next"]
    ring["next"] = (i + 1) % _HOST_RING
    if ring["bufs"][i] is None:
        ring["bufs"][i] = torch.empty(shape, dtype=dtype).pin_memory()
    if ring["ev"][i] is not None:
        ring["ev"][i].synchronize()       # (long complete in practice: two uploads ago)
    return ring, i


def _pack_threads():
    """Threads for the host pack: this process's fair share of the node's cores (torchrun exports LOCAL_WORLD_SIZE),
    never more than its CPU affinity allows, at most 16 — the pass is memory-bound."""
    import os
    total = os.cpu_count() or 1
    try:
        avail = len(os.sched_getaffinity(0))
    except Exception:
        avail = total
    ranks = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    return max(1, min(16, avail, max(1, total // ranks)))


def _pack_host_gt_u8(a):
    """float64 host crop (B,H,W,3) -> (pinned u8 tensor, ring, slot) if every value is an exact PNG code / 255.0
    (image_dataloaders.py:79-83), else None.  One threaded pass over the array (csrc/host_pack.cu)."""
    L = _lib.lib()
    ring, slot = _host_staging(a.shape, torch.uint8)
    buf = ring["bufs"][slot]
    ok = ctypes.c_int(0)
    rc = L.rac_host_pack_gt_u8(ctypes.c_void_p(a.__array_interface__["data"][0]), 1 if a.dtype == np.float64 else 0,
                               a.size, ctypes.c_void_p(buf.data_ptr()), _pack_threads(), ctypes.byref(ok))
    _lib.check(rc)
    return (buf, ring, slot) if ok.value else None


# ---- 8-bit gt crops ---------------------------------------------------------------------------------------------
# A uint8 (B,H,W,3) crop is taken as the raw `real.png` codes: the `/ 255.0` of image_dataloaders.py:79-83 moves into
# rac_gt_norm_u8, and the crop crosses PCIe at a quarter of the bytes of the f32 form.
def _is_u8(t):
    return (isinstance(t, torch.Tensor) and t.dtype == torch.uint8) or (isinstance(t, np.ndarray) and t.dtype == np.uint8)


def gt_norm_u8_device(gt_u8, x, use_norm_depth=False, _host_slot=None):
    """normalize_img(gt_u8 / 255.0) -> the three gt planes of x (B,C,H,W) f32 CUDA; gt_u8 (B,H,W,3) uint8, host or device.
    A pinned host tensor is uploaded on the library's copy stream (it overlaps whatever the current stream is running);
    the call returns only after that copy has read the caller's buffer (the reference's prepare_model_input is
    synchronous: a caller may refill its buffer right away).  _host_slot: (ring, slot) when gt_u8 is one of the
    library's own pinned staging buffers — the copy's event is parked on the slot instead of being waited for."""
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
        if _host_slot is not None:
            _host_slot[0]["ev"][_host_slot[1]] = ev
        else:
            ev.synchronize()
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


def _upload_f32_chunked(L, src, init, d, B, W, H, use_norm_depth, x, dev, host_convert):
    """gt crop `src` (host, (B,H,W,3)) -> normalised planes of x, in _GT_CHUNKS chunks of images: chunk k is copied on the
    library's copy stream while chunk k-1 is normalised + concatenated on the current stream (and the rasteriser of the
    call before is still running).  host_convert: src is pageable and/or not f32 — each chunk is first converted into a
    library-owned pinned buffer by a threaded torch copy, which overlaps the previous chunk's H2D copy."""
    main = torch.cuda.current_stream(dev)
    cs = _copy_stream(dev)
    ring, slot = _gt_staging(dev, (B, H, W, 3))
    gt = ring["bufs"][slot]
    if ring["done"][slot] is not None:
        cs.wait_event(ring["done"][slot])          # the staging buffer's previous consumer has finished
    if host_convert:
        hring, hslot = _host_staging((B, H, W, 3), torch.float32)
        hbuf = hring["bufs"][hslot]
        src_t = src if isinstance(src, torch.Tensor) else torch.from_numpy(src)
    per = (B + _GT_CHUNKS - 1) // _GT_CHUNKS
    events = []
    for b0 in range(0, B, per):
        b1 = min(B, b0 + per)
        if host_convert:
            hbuf[b0:b1].copy_(src_t[b0:b1])
            chunk = hbuf[b0:b1]
        else:
            chunk = src[b0:b1]
        with torch.cuda.stream(cs):
            gt[b0:b1].copy_(chunk, non_blocking=True)
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
    if host_convert:
        hring["ev"][hslot] = events[-1][2]
    else:
        # the caller's own pinned buffer: hand it back only once the copy stream has read all of it (the reference's
        # prepare_model_input is synchronous, so a drop-in caller may refill the buffer as soon as we return)
        events[-1][2].synchronize()


def assemble_input_device(init_hwc, gt_hwc, depth, use_norm_depth=False):
    """prepare_model_input on renders that already exist (iteration 0): -> x (B,C,H,W) f32 CUDA.
    gt_hwc, by what the caller holds (all paths normalise on the device):
      float64 host array (the unchanged loader, image_dataloaders.py:79-83) -> PNG codes recovered on the way into
          pinned staging (rac_host_pack_gt_u8), uploaded as bytes, rac_gt_norm_u8; float crops that are not exact
          codes / 255 fall through to the float path;
      uint8 (the PNG codes themselves) -> rac_gt_norm_u8;
      float32 / other host floats -> pinned staging (the caller's, or the library's after a threaded convert), chunked
          upload overlapped with normalisation;
      CUDA tensors -> used in place."""
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
    kv = _gt_cache_key(gt_hwc, dev) if _gt_cache_enabled else None      # (taken BEFORE anything is uploaded: it must describe
    cached = _gt_cache_lookup(gt_hwc, dev, kv)                          # the content the planes are computed from)
    if cached is not None:
        x[:, C - 3:].copy_(cached)
        _assemble_call(L, init, None, d, B, W, H, use_norm_depth, x, dev)
        return x
    capturing = torch.cuda.is_current_stream_capturing()
    if _is_u8(gt_hwc):
        gt_norm_u8_device(gt_hwc, x, use_norm_depth)
        _assemble_call(L, init, None, d, B, W, H, use_norm_depth, x, dev)
        _gt_cache_store(gt_hwc, dev, x[:, C - 3:], kv)
        return x
    host = _host_view(gt_hwc) if not capturing else None
    if host is not None and host.dtype == np.float64:
        packed = _pack_host_gt_u8(host)
        if packed is not None:
            buf, ring, slot = packed
            gt_norm_u8_device(buf, x, use_norm_depth, _host_slot=(ring, slot))
            _assemble_call(L, init, None, d, B, W, H, use_norm_depth, x, dev)
            _gt_cache_store(gt_hwc, dev, x[:, C - 3:], kv)
            return x
    if host is not None and host.dtype in (np.float32, np.float64) and B >= 2 * _GT_CHUNKS:
        _upload_f32_chunked(L, gt_hwc, init, d, B, W, H, use_norm_depth, x, dev, host_convert=not _overlappable(gt_hwc))
        _gt_cache_store(gt_hwc, dev, x[:, C - 3:], kv)
        return x
    gt = torch.as_tensor(gt_hwc).to(device=dev, dtype=torch.float32).contiguous()
    _assemble_call(L, init, gt, d, B, W, H, use_norm_depth, x, dev)
    _gt_cache_store(gt_hwc, dev, x[:, C - 3:], kv)
    return x


def to_numpy(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
