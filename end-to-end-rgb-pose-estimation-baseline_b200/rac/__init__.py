"""rac — the B200-native render-and-compare core (ctypes over librac_b200.so).

Public surface used by the shadow modules one directory up (renderer.py, data_loaders.py,
rotation_representation.py, loss.py, se3_helpers.py) and by callers that want to stay on the
device: rac.render_batch_device, rac.render_concat_device, rac.assemble_input_device,
rac.mesh (load / procedural / device cache), rac.ops (autograd functions), and the two rows next to the path
(SURVEY.md §8f): rac.sample_poses_device, rac.silhouette_device, rac.blend_device.
"""
from . import _lib, mesh, ops
from ._lib import RacError, launch_count, lib
from .aux import blend_device, sample_poses_device, silhouette_device
from .render import assemble_input_device, gt_norm_u8_device, render_batch_device, render_concat_device
from .step import RenderCompareStep

__all__ = ["_lib", "mesh", "ops", "RacError", "launch_count", "lib", "assemble_input_device",
           "render_batch_device", "render_concat_device", "gt_norm_u8_device", "RenderCompareStep", "sample_poses_device", "silhouette_device",
           "blend_device"]
