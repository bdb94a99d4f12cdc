"""ctypes binding of librac_b200.so (include/rac.h).  No fallback: if the shared library
cannot be loaded, or no CUDA device is present when a kernel is needed, this raises."""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
# RAC_LIB selects a diagnostic build of the same sources (profiles/phase_timing.py); the default is the product library
LIB_PATH = os.environ.get("RAC_LIB") or os.path.join(_PKG, "librac_b200.so")

c_int = ctypes.c_int
c_void_p = ctypes.c_void_p
c_size_t = ctypes.c_size_t
c_float = ctypes.c_float
c_double = ctypes.c_double
c_int_p = ctypes.POINTER(ctypes.c_int)


class RacMaterial(ctypes.Structure):
    _fields_ = [("base_color", c_float * 4), ("metallic", c_float), ("roughness", c_float)]


class RacPart(ctypes.Structure):
    _fields_ = [("material", RacMaterial), ("tex_rgb8", c_void_p), ("tex_w", c_int), ("tex_h", c_int),
                ("use_vertex_color", c_int), ("double_sided", c_int)]


class RacScene(ctypes.Structure):
    _fields_ = [("ambient", c_float), ("light_intensity", c_float), ("inner_cone", c_float),
                ("outer_cone", c_float), ("znear", c_float), ("zfar", c_float)]


# every symbol include/rac.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "rac_last_error": (ctypes.c_char_p, []),
    "rac_version": (c_int, []),
    "rac_device_info": (c_int, [c_int_p, c_int_p, c_int_p]),
    "rac_launch_count": (ctypes.c_longlong, []),
    "rac_scene_defaults": (None, [ctypes.POINTER(RacScene)]),
    "rac_material_defaults": (None, [ctypes.POINTER(RacMaterial)]),
    "rac_mesh_create": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p,
                                ctypes.POINTER(RacMaterial), c_int_p]),
    "rac_mesh_create_parts": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int_p]),
    "rac_mesh_destroy": (c_int, [c_int]),
    "rac_mesh_info": (c_int, [c_int, c_int_p, c_int_p]),
    "rac_render_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "rac_render_batch": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, ctypes.POINTER(RacScene),
                                 c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "rac_render_concat": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, ctypes.POINTER(RacScene),
                                  c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "rac_assemble_workspace_bytes": (c_size_t, [c_int]),
    "rac_assemble_input": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                   c_size_t, c_void_p]),
    "rac_assemble_input_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "rac_pose_update_fwd": (c_int, [c_void_p, c_int, c_void_p, c_int, c_double, c_double, c_int, c_void_p, c_void_p,
                                    c_void_p]),
    "rac_pose_update_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "rac_rotation_fwd": (c_int, [c_void_p, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p]),
    "rac_rotation_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "rac_add_loss_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_void_p, c_void_p]),
    "rac_add_loss_bwd": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "rac_pose_loss_step": (c_int, [c_void_p, c_int, c_void_p, c_int, ctypes.c_double, ctypes.c_double, c_void_p, c_void_p, c_void_p,
                                   c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "rac_transform_pts_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "rac_transform_pts_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "rac_sample_poses": (c_int, [c_void_p, c_void_p, c_int, c_double, c_double, c_double, c_double, c_double,
                                 c_void_p, c_void_p, c_void_p]),
    "rac_silhouette": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "rac_blend": (c_int, [c_void_p, c_void_p, c_float, c_int, c_size_t, c_void_p, c_void_p]),
    "rac_gt_u8_workspace_bytes": (c_size_t, [c_int]),
    "rac_gt_norm_u8": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "rac_host_pack_gt_u8": (c_int, [c_void_p, c_int, c_size_t, c_void_p, c_int, ctypes.POINTER(c_int)]),
}

RAC_ROT = {"6D": 0, "SVD": 1, "3D": 2}
RAC_POSE_SAVED_FLOATS = 48
LOSS_ADD_L1, LOSS_ADD_L2, LOSS_DISENTANGLED, LOSS_DISENT_SCALED = 0, 1, 2, 3

_lib = None
_lock = threading.Lock()


class RacError(RuntimeError):
    pass


def lib():
    """Returns the loaded library; builds it in-tree with nvcc if the .so is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            import importlib.util
            spec = importlib.util.spec_from_file_location("_rac_build", os.path.join(_PKG, "build.py"))
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build()
        try:
            handle = ctypes.CDLL(LIB_PATH)
        except OSError as e:
            raise ImportError(f"librac_b200.so could not be loaded from {LIB_PATH}: {e}. "
                              "There is no CPU or PyTorch fallback for the render-and-compare kernels.") from e
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
        return _lib


def last_error():
    return lib().rac_last_error().decode("utf-8", "replace")


def check(rc, exc=RacError):
    """rc < 0 -> raise.  RAC_E_INVALID surfaces as AssertionError (the reference asserts on bad shapes)."""
    if rc == 0:
        return
    msg = last_error()
    if rc == -1:
        raise AssertionError(msg)
    raise exc(f"librac_b200 error {rc}: {msg}")


_cuda_ok = False


def require_cuda():
    global _cuda_ok
    import torch
    if not _cuda_ok:          # (checked until it succeeds once: a process does not lose its CUDA device)
        if not torch.cuda.is_available():
            raise RacError("no CUDA device: the render-and-compare path is CUDA-only (sm_100a) and has no CPU fallback")
        _cuda_ok = True
    return torch.device("cuda", torch.cuda.current_device())


_raw_stream = None


def raw_stream(device_index=None):
    """cudaStream_t of torch's current stream as an int — torch._C._cuda_getCurrentRawStream where it exists (one C call;
    torch.cuda.current_stream() builds a Stream object through several Python layers, ~10 us, and every C-ABI call of
    the hot path needs the handle)."""
    global _raw_stream
    import torch
    if _raw_stream is None:
        fast = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        _raw_stream = fast if fast is not None else (lambda i: torch.cuda.current_stream(i).cuda_stream)
    return _raw_stream(torch.cuda.current_device() if device_index is None else device_index)


def stream_ptr():
    return c_void_p(raw_stream())


def ptr(t):
    """device/host pointer of a tensor (or None)."""
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


def launch_count():
    return int(lib().rac_launch_count())
