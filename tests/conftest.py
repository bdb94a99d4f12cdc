import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (PKG, ROOT, GOLDEN):
    if p not in sys.path:
        sys.path.insert(0, p)


collect_ignore = ["ref_drivers"]      # the reference's own test_model.py lives there as a fixture, not as a test


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


SCENE_CFG = dict(distance_cam_to_world=2.5, distance_cam_to_world_deviation=0.1,
                 world_to_object_gt_transl_deviation=0.1, world_to_object_transl_deviation=0.1,
                 world_to_object_angle_deviation=30)


def sample_poses(B, seed):
    """(T_init, T_gt) f32 (B,4,4) drawn exactly like se3_helpers.get_T_CO_init_and_gt with the
    example config's scene parameters (configs/example_config.py:45-52)."""
    import se3_helpers
    state = np.random.get_state()
    np.random.seed(seed)
    ti, tg = [], []
    for _ in range(B):
        a, b = se3_helpers.get_T_CO_init_and_gt(SCENE_CFG)
        ti.append(a.data[0])
        tg.append(b.data[0])
    np.random.set_state(state)
    return np.array(ti, np.float32), np.array(tg, np.float32)


def intrinsics(B, W, H, focal=50.0, sensor=36.0):
    """renderer.get_camera_matrix with 50 mm / 36 mm (old-configs/effnetv2_m_svd.py:49-51) at width W."""
    fx = focal / (sensor / W)
    K = np.array([[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]], np.float32)
    return np.tile(K, (B, 1, 1))


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
