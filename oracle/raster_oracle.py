"""ctypes front-end of oracle/raster_ref.c (CPU oracle of steps 1-2).  TEST INFRASTRUCTURE ONLY.

Parity status: UNPINNED — see the header of raster_ref.c.
"""
import ctypes
import os

import numpy as np

from . import build as _build

_lib = None
DEFAULT_SCENE = np.array([0.1, 15.0, np.pi / 9.0, np.pi / 2.0, 0.05, 100.0], dtype=np.float32)
DEFAULT_MATERIAL = np.array([0.3, 0.3, 0.3, 1.0, 0.2, 0.8], dtype=np.float32)


def lib():
    global _lib
    if _lib is None:
        path = _build.build()
        L = ctypes.CDLL(path)
        vp, ci = ctypes.c_void_p, ctypes.c_int
        L.oracle_render_batch.restype = ci
        L.oracle_render_batch.argtypes = [vp, ci, vp, ci, vp, vp, vp, ci, ci, vp, vp, vp, vp, ci, ci, ci, vp, vp, vp, ci]
        L.oracle_render_batch_parts.restype = ci
        L.oracle_render_batch_parts.argtypes = [vp, ci, vp, ci, vp, vp, vp, ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ci, ci, vp, vp, vp, ci]
        L.oracle_prepare_model_input.restype = ci
        L.oracle_prepare_model_input.argtypes = [vp, vp, vp, ci, ci, ci, ci, vp, ci]
        L.oracle_num_threads.restype = ci
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def render_batch(mesh, T_CO, K, img_size, scene=None, nthreads=0, want_rgba8=True):
    """mesh: object with verts (V,3), faces (T,3) and optional uv / vcolor / tex / material_array().
    T_CO (B,4,4), K (B,3,3).  img_size: int or (W,H).
    -> rgb (B,H,W,3) f32 max-normalised, depth (B,H,W) f32, rgba8 (B,H,W,4) u8."""
    W, H = (img_size, img_size) if np.isscalar(img_size) else (int(img_size[0]), int(img_size[1]))
    T = np.ascontiguousarray(T_CO, dtype=np.float32)
    Kc = np.ascontiguousarray(K, dtype=np.float32)
    B = T.shape[0]
    assert T.shape == (B, 4, 4) and Kc.shape == (B, 3, 3)
    verts = np.ascontiguousarray(mesh.verts, dtype=np.float32)
    faces = np.ascontiguousarray(mesh.faces, dtype=np.int32)
    uv = None if getattr(mesh, "uv", None) is None else np.ascontiguousarray(mesh.uv, dtype=np.float32)
    vc = None if getattr(mesh, "vcolor", None) is None else np.ascontiguousarray(mesh.vcolor, dtype=np.uint8)
    tex = None if getattr(mesh, "tex", None) is None else np.ascontiguousarray(mesh.tex, dtype=np.uint8)
    th, tw = (tex.shape[0], tex.shape[1]) if tex is not None else (0, 0)
    mat = np.ascontiguousarray(mesh.material_array() if hasattr(mesh, "material_array") else DEFAULT_MATERIAL, dtype=np.float32)
    sc = np.ascontiguousarray(DEFAULT_SCENE if scene is None else scene, dtype=np.float32)
    rgb = np.empty((B, H, W, 3), dtype=np.float32)
    depth = np.empty((B, H, W), dtype=np.float32)
    rgba8 = np.empty((B, H, W, 4), dtype=np.uint8) if want_rgba8 else None
    parts = getattr(mesh, "parts", None)
    if parts:       # multi-part mesh (GLB scene): material / texture per face
        n = len(parts)
        pmat = np.ascontiguousarray([np.asarray(p["material"], np.float32).reshape(6) for p in parts], dtype=np.float32)
        texs = [None if (p.get("tex") is None or uv is None) else np.ascontiguousarray(p["tex"], dtype=np.uint8) for p in parts]
        ptex = (ctypes.c_void_p * n)(*[t.ctypes.data if t is not None else None for t in texs])
        ptw = np.ascontiguousarray([t.shape[1] if t is not None else 0 for t in texs], dtype=np.int32)
        pth = np.ascontiguousarray([t.shape[0] if t is not None else 0 for t in texs], dtype=np.int32)
        puse = np.ascontiguousarray([int(bool(p.get("use_vcolor", False))) for p in parts], dtype=np.int32)
        pds = np.ascontiguousarray([int(bool(p.get("double_sided", False))) for p in parts], dtype=np.int32)
        fp = np.ascontiguousarray(mesh.face_part, dtype=np.uint8)
        rc = lib().oracle_render_batch_parts(_p(verts), len(verts), _p(faces), len(faces), _p(uv), _p(vc), _p(fp), n, _p(pmat),
                                             ctypes.cast(ptex, ctypes.c_void_p), _p(ptw), _p(pth), _p(puse), _p(pds), _p(sc), _p(T), _p(Kc),
                                             B, W, H, _p(rgb), _p(depth), _p(rgba8), int(nthreads))
    else:
        rc = lib().oracle_render_batch(_p(verts), len(verts), _p(faces), len(faces), _p(uv), _p(vc), _p(tex), tw, th,
                                       _p(mat), _p(sc), _p(T), _p(Kc), B, W, H, _p(rgb), _p(depth), _p(rgba8), int(nthreads))
    if rc != 0:
        raise MemoryError("oracle_render_batch failed")
    return rgb, depth, rgba8


def render_batch_mixed(meshes, T_CO, K, img_size, scene=None, nthreads=0):
    """Per-image meshes (list of B): groups by mesh object and renders each group."""
    B = len(meshes)
    W, H = (img_size, img_size) if np.isscalar(img_size) else (int(img_size[0]), int(img_size[1]))
    rgb = np.empty((B, H, W, 3), np.float32)
    depth = np.empty((B, H, W), np.float32)
    rgba8 = np.empty((B, H, W, 4), np.uint8)
    groups = {}
    for i, m in enumerate(meshes):
        groups.setdefault(id(m), (m, []))[1].append(i)
    for m, idx in groups.values():
        r, d, q = render_batch(m, np.asarray(T_CO)[idx], np.asarray(K)[idx], img_size, scene, nthreads)
        rgb[idx], depth[idx], rgba8[idx] = r, d, q
    return rgb, depth, rgba8


def prepare_model_input_c(init, gt, depth, use_norm_depth=False, nthreads=0):
    """C restatement of data_loaders.py:158-177 (the timed CPU baseline; checked against
    oracle/assemble_ref.py in tests)."""
    init = np.ascontiguousarray(init, dtype=np.float32)
    gt = np.ascontiguousarray(gt, dtype=np.float32)
    B, H, W, _ = init.shape
    d = None if depth is None else np.ascontiguousarray(depth, dtype=np.float32)
    x = np.empty((B, 7 if use_norm_depth else 6, H, W), dtype=np.float32)
    lib().oracle_prepare_model_input(_p(init), _p(gt), _p(d), B, W, H, int(bool(use_norm_depth)), _p(x), int(nthreads))
    return x


def num_threads():
    return int(lib().oracle_num_threads())
