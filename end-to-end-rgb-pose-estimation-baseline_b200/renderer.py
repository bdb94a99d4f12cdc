"""Drop-in for the reference's renderer.py (the module train/test import as `renderer`).

Same public names and signatures (ref: renderer.py:17-25 get_camera_matrix, :85-105
render_scene); the pyrender/OpenGL scene is replaced by the sm_100a rasteriser in
csrc/raster.cu.  The mesh is parsed and uploaded once per path instead of once per image
(ref: renderer.py:99 -> :28-48).  `render_normals` / `plot_SE3` are dead code in the
reference (missing shader assets; plotting) and are not provided.
"""
import numpy as np

import rac
from rac.lazy import LazyArray   # noqa: F401  (re-export for callers that test for it)


def get_camera_matrix(intrinsics):
    """ref: renderer.py:17-25.  fx = fy = focal_length / (sensor_width / image_resolution)."""
    f = intrinsics["focal_length"]
    res = intrinsics["image_resolution"]
    mm_per_px = intrinsics["sensor_width"] / res
    fx = fy = f / mm_per_px
    c = res / 2
    return np.array([[fx, 0, c], [0, fy, c], [0, 0, 1]])


def render_scene(object_path, T_CO, cam_config=None, K=None, img_size=None):
    """ref: renderer.py:85-105.  -> (img (S,S,3) float64 in [0,1], max-normalised; depth (S,S) f32)."""
    T_CO = np.asarray(T_CO)
    assert T_CO.shape == (4, 4)
    if not (K is not None and img_size is not None):
        assert cam_config is not None
        K = get_camera_matrix(cam_config)
        img_size = cam_config["image_resolution"]
    K = np.asarray(K.detach().cpu() if hasattr(K, "detach") else K, dtype=np.float64)
    out = rac.render_batch_device(T_CO[None].astype(np.float32), [object_path], K[None].astype(np.float32), img_size)
    img = out["rgb"][0].cpu().numpy().astype(np.float64)
    depth = out["depth"][0].cpu().numpy()
    return img, depth
