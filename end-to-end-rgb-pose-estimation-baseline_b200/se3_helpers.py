"""Drop-in for the reference's se3_helpers.py (pose sampling for dataset generation; host-side
numpy, not on the train/test hot path — SURVEY.md F11).  Same names and signatures
(ref: se3_helpers.py:5-88); spatialmath is replaced by rac.sm_lite (same `.data[0]`, `*`, `.inv()`)."""
import numpy as np

from rac import sm_lite as sm


def look_at_SE3(origin, target, up):
    """ref: se3_helpers.py:5-25.  Camera pose (camera z towards target)."""
    assert len(origin) == 3
    assert len(target) == 3
    assert len(up) == 3
    origin, target, up = np.array(origin), np.array(target), np.array(up)
    z = target - origin
    z = z / np.linalg.norm(z)
    x = np.cross(up, z)
    x = -x / np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.array([x, y, z]).transpose()
    return sm.SE3.Rt(sm.SO3(R, check=True), origin)


def get_random_z_rot():
    return sm.SE3.Rz(np.random.uniform(0.0, 180.0), unit="deg")


def get_random_unit_axis():
    """ref: se3_helpers.py:32-37 — uniform in a cube, then normalised (deliberately not uniform on the sphere)."""
    a = np.random.random(3) - 0.5
    return a / np.linalg.norm(a)


def apply_small_random_rotation_translation(T, theta_range_deg, xyz_transl_range, sampling="uniform"):
    """ref: se3_helpers.py:39-57 — delta * T (left-multiplied perturbation)."""
    if sampling == "uniform":
        theta = np.random.uniform(-theta_range_deg, theta_range_deg)
        tx = np.random.uniform(-xyz_transl_range, xyz_transl_range)
        ty = np.random.uniform(-xyz_transl_range, xyz_transl_range)
        tz = np.random.uniform(-xyz_transl_range, xyz_transl_range)
    if sampling == "normal":
        theta = np.random.normal(0, theta_range_deg)
        tx = np.random.normal(0, xyz_transl_range)
        ty = np.random.normal(0, xyz_transl_range)
        tz = np.random.normal(0, xyz_transl_range)
    axis = get_random_unit_axis()
    delta = sm.SE3.Rt(sm.SO3.AngleAxis(theta, axis, unit="deg"), np.array([tx, ty, tz]))
    return delta * T


def get_random_rotation_translation(xyz_transl_range):
    """ref: se3_helpers.py:59-68."""
    theta = np.random.uniform(0, 360)
    axis = get_random_unit_axis()
    R = sm.SO3.AngleAxis(theta, axis, unit="deg")
    t = np.array([np.random.normal(0, xyz_transl_range) for _ in range(3)])
    return sm.SE3.Rt(R, t)


def get_T_CW(base_distance, random_deviation=0.0):
    """ref: se3_helpers.py:70-74 — camera on +x looking at the origin, up +z."""
    d = base_distance + np.random.uniform(-random_deviation, random_deviation)
    return look_at_SE3([d, 0, 0], [0, 0, 0], [0, 0, 1]).inv()


def get_T_CO_init_and_gt(scene_config):
    """ref: se3_helpers.py:76-88."""
    T_WO_gt = get_random_rotation_translation(scene_config["world_to_object_gt_transl_deviation"])
    T_WO_init = apply_small_random_rotation_translation(T_WO_gt, scene_config["world_to_object_angle_deviation"],
                                                        scene_config["world_to_object_transl_deviation"])
    T_CW = get_T_CW(scene_config["distance_cam_to_world"], scene_config["distance_cam_to_world_deviation"])
    return T_CW * T_WO_init, T_CW * T_WO_gt
