"""Drop-in for the hot-path half of the reference's data_loaders.py.

Same public names and signatures: render_batch / render_batch_sequential (ref:
data_loaders.py:84-118), normalize_depth (:69-77), normalize_img (:79-82),
prepare_model_input (:158-177), get_camera_mat_tensor (:181-191), and the vertex / pose
sampling helpers the drivers pull in through `from data_loaders import *` (:19-67,148-154).
Rendering and input assembly run as CUDA kernels (csrc/raster.cu); there is no CPU path.
The broken thread-per-image variant (:120-146, unreachable in the reference) is not provided:
parallel_render=True renders the same batch with the same kernels.
"""
import os
import random
import time

import numpy as np
import torch

import rac
from rac.lazy import LazyArray
from rac.mesh import host_mesh
from renderer import render_scene   # noqa: F401  (ref: data_loaders.py:11)
from se3_helpers import get_T_CO_init_and_gt


def _np(x):
    return x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)


# ---- mesh / pose sampling helpers (host side, ref: data_loaders.py:19-67) ---------------------
def get_vertices(mesh_path):
    return host_mesh(mesh_path).vertices


def sample_vertices_as_tensor(mesh_path, num_verts=1000):
    """ref: data_loaders.py:24-32 — num_verts vertices drawn with replacement via `random.choice`."""
    verts = get_vertices(mesh_path)
    picked = [np.array(random.choice(verts)) for _ in range(num_verts)]
    return torch.tensor(np.array(picked, dtype=np.float32))


def get_dataset_class_paths(dataset_name, classname, test_or_train):
    d = os.path.join("model3d-datasets", dataset_name, classname, test_or_train)
    return [os.path.join(d, name) for name in os.listdir(d)]


def get_all_train_or_test_paths(ds_name, classes, train_or_test):
    paths = []
    for cls in classes:
        paths = paths + get_dataset_class_paths(ds_name, cls, train_or_test)
    return paths


def sample_random_path(ds_name, classes, train_or_test):
    return random.choice(get_all_train_or_test_paths(ds_name, classes, train_or_test))


def sample_mesh_paths(batch_size, ds_name, classes, train_or_test):
    return [sample_random_path(ds_name, classes, train_or_test) for _ in range(batch_size)]


def sample_T_CO_inits_and_gts(batch_size, scene_config):
    inits, gts = [], []
    for _ in range(batch_size):
        T_init, T_gt = get_T_CO_init_and_gt(scene_config)
        inits.append(T_init.data[0].astype(np.float32))
        gts.append(T_gt.data[0].astype(np.float32))
    return np.array(inits, dtype=np.float32), np.array(gts, dtype=np.float32)


def sample_T_CO_inits_and_gts_device(batch_size, scene_config, generator=None):
    """The same distribution as sample_T_CO_inits_and_gts, composed for the whole batch by one kernel
    (rac_sample_poses) from torch's device RNG: -> two (B,4,4) f32 CUDA tensors."""
    return rac.sample_poses_device(batch_size, scene_config, generator=generator)


def sample_verts_to_batch(mesh_paths, num_verts_to_sample):
    return torch.stack([sample_vertices_as_tensor(p, num_verts_to_sample) for p in mesh_paths])


def get_camera_mat_tensor(cam_intrinsics, batch_size):
    """ref: data_loaders.py:181-191."""
    res = cam_intrinsics["image_resolution"]
    px = cam_intrinsics["sensor_width"] / res
    fx = fy = cam_intrinsics["focal_length"] / px
    c = res / 2
    K = torch.tensor(np.array([[fx, 0.0, c], [0.0, fy, c], [0.0, 0.0, 1.0]], dtype=np.float32))
    return K.unsqueeze(0).repeat(batch_size, 1, 1)


# ---- step 1: render ------------------------------------------------------------------------
def render_batch(T_COs, mesh_paths, Ks, img_size, parallel_render=False):
    """ref: data_loaders.py:84-98.  -> (imgs (B,S,S,3) f32, depths (B,S,S) f32) as lazily
    materialised host arrays (rac.lazy.LazyArray; np.asarray(x) gives the ndarray)."""
    t0 = time.time()
    imgs, depths = render_batch_sequential(T_COs, mesh_paths, Ks, img_size)
    print("Par render_time:" if parallel_render else "Seq render_time:", time.time() - t0)
    return imgs, depths


def render_batch_sequential(T_COs, mesh_paths, Ks, img_size):
    """ref: data_loaders.py:100-118 — one kernel pipeline for the whole batch instead of a Python loop."""
    bsz = len(mesh_paths)
    T = T_COs.device_tensor if isinstance(T_COs, LazyArray) else T_COs
    assert tuple(T.shape) == (bsz, 4, 4)
    assert tuple(Ks.shape) == (bsz, 3, 3)
    out = rac.render_batch_device(T, list(mesh_paths), Ks, img_size)
    return LazyArray(out["rgb"]), LazyArray(out["depth"])


# ---- step 2: input assembly -------------------------------------------------------------------
def normalize_depth(depth_img):
    """ref: data_loaders.py:69-77 (single image).  Runs the assembly kernel's depth branch."""
    d = _np(depth_img).astype(np.float32)
    H, W = d.shape
    z = np.zeros((1, H, W, 3), dtype=np.float32)
    z[0, 0, 0, 0] = 1.0   # any non-constant gt keeps the (unused) gt channels finite
    x = rac.assemble_input_device(z, z, d[None], use_norm_depth=True)
    return x[0, 0].cpu().numpy()


def normalize_img(img):
    """ref: data_loaders.py:79-82 (single image): (img - mean) / std over all values, ddof 0."""
    a = _np(img)
    H, W, _ = a.shape
    x = rac.assemble_input_device(np.zeros((1, H, W, 3), np.float32), a[None].astype(np.float32), None, False)
    return x[0, 3:6].permute(1, 2, 0).cpu().numpy().astype(a.dtype if a.dtype.kind == "f" else np.float32)


def prepare_model_input(init_imgs, gt_imgs, depths, use_norm_depth=False):
    """ref: data_loaders.py:158-177.  -> (B,C,S,S) f32 tensor, channels [norm_depth?, render RGB,
    normalised gt RGB].  The tensor lives on the GPU (the reference builds it on the host and the
    caller moves it with `.to(device)`, train_model.py:243 — that call is now a no-op)."""
    dev = lambda a: a.device_tensor if isinstance(a, LazyArray) else a
    return rac.assemble_input_device(dev(init_imgs), dev(gt_imgs), dev(depths) if use_norm_depth else None,
                                     use_norm_depth)
