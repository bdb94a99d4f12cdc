"""Generates tests/golden/aux_golden.npz — golden vectors for SURVEY.md §8f N3/N4, made in the build container with
the libraries the reference itself calls:
  silhouette  scipy.signal.convolve2d, exactly as /root/reference/visualization.py:19-31
  blend       PIL.Image.blend,        exactly as /root/reference/visualization.py:34-37
  poses       this repo's se3_helpers.get_T_CO_init_and_gt (the shadow of /root/reference/se3_helpers.py:76-88 over
              rac.sm_lite; the reference module itself needs spatialmath, which is not installed) with numpy's global
              RNG replaced by a recorded stream, so the variates are known.
Run: python tests/golden/make_aux_golden.py"""
import os
import sys

import numpy as np
from PIL import Image
from scipy.signal import convolve2d

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200"), ROOT]

CFG = dict(distance_cam_to_world=2.5, distance_cam_to_world_deviation=0.1, world_to_object_gt_transl_deviation=0.1,
           world_to_object_transl_deviation=0.1, world_to_object_angle_deviation=30)


def ref_silhouette(img, only):
    img = img.copy()
    gray = np.mean(img, axis=2)
    sil = np.where(gray > 0, 1.0, 0.0)
    res = np.clip(convolve2d(sil, np.ones((5, 5)), mode="same"), 0.0, 1.0)
    sil = res - sil
    if only:
        return sil
    img[sil > 0] = (1.0, 0.0, 0.0)
    return img


def ref_blend(a, b, alpha):
    return np.asarray(Image.blend(Image.fromarray(np.uint8(a * 255.0)), Image.fromarray(np.uint8(b * 255.0)), alpha)) / 255.0


class RecordedRng:
    """Stands in for np.random.{uniform,normal,random}: serves U(0,1) / N(0,1) draws from fixed arrays and maps them
    the way numpy does (low + (high - low) u; loc + scale n), recording nothing else."""

    def __init__(self, u, n):
        self.u, self.n, self.iu, self.inn = u, n, 0, 0

    def _u(self):
        v = self.u[self.iu]; self.iu += 1
        return float(v)

    def uniform(self, low=0.0, high=1.0):
        return low + (high - low) * self._u()

    def random(self, size):
        return np.array([self._u() for _ in range(size)])

    def normal(self, loc=0.0, scale=1.0):
        v = self.n[self.inn]; self.inn += 1
        return loc + scale * float(v)


def ref_poses(u01, n01):
    import se3_helpers
    Ti, Tg = [], []
    for b in range(len(u01)):
        # draw order of the reference (se3_helpers.py:59-68, 39-57, 70-74):
        # theta_gt, axis_gt(3) | n n n | theta_d, tx, ty, tz, axis_d(3) | distance
        rr = RecordedRng(u01[b], n01[b])
        saved = (np.random.uniform, np.random.random, np.random.normal)
        np.random.uniform, np.random.random, np.random.normal = rr.uniform, rr.random, rr.normal
        try:
            a, g = se3_helpers.get_T_CO_init_and_gt(CFG)
        finally:
            np.random.uniform, np.random.random, np.random.normal = saved
        assert rr.iu == 12 and rr.inn == 3
        Ti.append(a.data[0]); Tg.append(g.data[0])
    return np.array(Ti, np.float32), np.array(Tg, np.float32)


def main():
    rng = np.random.default_rng(42)
    H, W = 48, 64
    img = np.zeros((H, W, 3), np.float32)
    img[10:30, 14:40] = rng.uniform(0.05, 1, (20, 26, 3))
    img[0:3, 0:5] = 0.5; img[45:, 60:] = 0.25; img[20, 50] = 0.7
    a = rng.uniform(0, 1, (H, W, 3)).astype(np.float32)
    b64 = rng.integers(0, 256, (H, W, 3)) / 255.0
    u01 = rng.uniform(0, 1, (16, 12)).astype(np.float32)
    n01 = rng.normal(size=(16, 3)).astype(np.float32)
    Ti, Tg = ref_poses(u01.astype(np.float64), n01.astype(np.float64))
    out = dict(img=img, sil_only=ref_silhouette(img, True), sil_paint=ref_silhouette(img, False),
               blend_a=a, blend_b64=b64, blend_f32_03=ref_blend(a, b64.astype(np.float32), 0.3),
               blend_mixed_05=ref_blend(a, b64, 0.5), blend_f64_077=ref_blend(a.astype(np.float64), b64, 0.77),
               u01=u01, n01=n01, T_init=Ti, T_gt=Tg)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "aux_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
