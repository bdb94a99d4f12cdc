"""CPU ORACLE (test infrastructure, never shipped or measured as the product) for the two rows next to the hot path
(SURVEY.md §8f): numpy restatements of
  create_silhouette / blend_imgs   — /root/reference/visualization.py:19-37
  get_T_CO_init_and_gt             — /root/reference/se3_helpers.py:5-88, driven by explicit variates
Pinned: tests/test_oracle_pinned.py checks silhouette() against scipy.signal.convolve2d (what the reference calls)
and blend() against PIL.Image.blend (ditto); tests/golden/aux_golden.npz holds outputs generated that way."""
import numpy as np


def silhouette(img, silhouette_only=False):
    """visualization.py:19-31.  Returns a new array (the reference paints in place and returns its argument)."""
    gray = np.mean(img, axis=2)
    sil = np.where(gray > 0, 1.0, 0.0)
    H, W = sil.shape
    pad = np.zeros((H + 4, W + 4))
    pad[2:-2, 2:-2] = sil
    res = np.zeros((H, W))
    for dy in range(5):                       # convolve2d(sil, ones(5,5), mode='same'): zero padded box sum
        for dx in range(5):
            res += pad[dy:dy + H, dx:dx + W]
    ring = np.clip(res, 0.0, 1.0) - sil
    if silhouette_only:
        return ring
    out = np.array(img, copy=True)
    out[ring > 0] = (1.0, 0.0, 0.0)
    return out


def blend(im1, im2, alpha):
    """visualization.py:34-37 with PIL's Image.blend restated: per channel (uint8)(a + alpha * (b - a)), the sum
    formed in C float (f32), truncated."""
    a = np.uint8(im1 * 255.0).astype(np.float32)
    b = np.uint8(im2 * 255.0).astype(np.float32)
    t = a + np.float32(alpha) * (b - a)
    return np.uint8(t).astype(np.float64) / 255.0


def _angle_axis(theta_deg, v):
    th = np.deg2rad(theta_deg)
    n = np.linalg.norm(v)
    eps = 10 * np.finfo(np.float64).eps
    if n < eps or abs(th) < eps:
        return np.eye(3)
    k = v / n
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def _look_at(origin, target, up):
    """se3_helpers.py:5-22."""
    origin, target, up = (np.asarray(a, dtype=np.float64) for a in (origin, target, up))
    z = target - origin
    z = z / np.linalg.norm(z)
    x = np.cross(up, z)
    x = -x / np.linalg.norm(x)
    y = np.cross(z, x)
    T = np.eye(4)
    T[:3, :3] = np.array([x, y, z]).transpose()
    T[:3, 3] = origin
    return T


def poses_from_variates(u01, n01, cfg):
    """se3_helpers.py:76-88 for a batch, with the random draws made explicit: u01 (B,12) uniforms in [0,1) in the
    order [theta_gt, axis_gt(3), theta_delta, t_delta(3), axis_delta(3), cam distance], n01 (B,3) standard normals
    (gt translation).  -> (T_CO_init, T_CO_gt) f32 (B,4,4)."""
    u01 = np.asarray(u01, dtype=np.float64)
    n01 = np.asarray(n01, dtype=np.float64)
    B = len(u01)
    Ti, Tg = np.zeros((B, 4, 4), np.float32), np.zeros((B, 4, 4), np.float32)
    a, r = cfg["world_to_object_angle_deviation"], cfg["world_to_object_transl_deviation"]
    for b in range(B):
        u, g = u01[b], n01[b]
        T_WO_gt = np.eye(4)                                                   # :59-68
        T_WO_gt[:3, :3] = _angle_axis(360.0 * u[0], u[1:4] - 0.5)
        T_WO_gt[:3, 3] = cfg["world_to_object_gt_transl_deviation"] * g
        delta = np.eye(4)                                                     # :39-57
        delta[:3, :3] = _angle_axis(-a + 2 * a * u[4], u[8:11] - 0.5)
        delta[:3, 3] = -r + 2 * r * u[5:8]
        T_WO_init = delta @ T_WO_gt
        d = cfg["distance_cam_to_world"] + (-1 + 2 * u[11]) * cfg["distance_cam_to_world_deviation"]   # :70-74
        T_CW = np.linalg.inv(_look_at([d, 0, 0], [0, 0, 0], [0, 0, 1]))
        Ti[b] = (T_CW @ T_WO_init).astype(np.float32)
        Tg[b] = (T_CW @ T_WO_gt).astype(np.float32)
    return Ti, Tg
