"""GPU parity for the rows next to the hot path, through the C ABI (rac_sample_poses,
rac_silhouette, rac_blend) and the shadow module visualization.py, against oracle/aux_ref.py and the golden vectors."""
import os

import numpy as np
import pytest
import torch

from conftest import SCENE_CFG
from oracle import aux_ref

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "aux_golden.npz")


@pytest.fixture(scope="module")
def rac():
    import rac as _rac
    _rac.lib()
    return _rac


@pytest.fixture(scope="module")
def gold():
    return {k: v for k, v in np.load(GOLD).items()}


def test_sample_poses_matches_oracle_and_golden(rac, gold):
    Ti, Tg = rac.sample_poses_device(16, SCENE_CFG, variates=(gold["u01"], gold["n01"]))
    # tolerance: f64 composition on both sides, one f32 rounding at the end (1e-5 relative is the north-star bar)
    np.testing.assert_allclose(Ti.cpu().numpy(), gold["T_init"], rtol=1e-6, atol=3e-7)
    np.testing.assert_allclose(Tg.cpu().numpy(), gold["T_gt"], rtol=1e-6, atol=3e-7)
    rng = np.random.default_rng(3)
    B = 1000
    u, n = rng.uniform(0, 1, (B, 12)).astype(np.float32), rng.normal(size=(B, 3)).astype(np.float32)
    u[0, 4] = 0.5                       # zero perturbation angle -> identity rotation branch
    Ti, Tg = rac.sample_poses_device(B, SCENE_CFG, variates=(u, n))
    Ri, Rg = aux_ref.poses_from_variates(u, n, SCENE_CFG)
    np.testing.assert_allclose(Ti.cpu().numpy(), Ri, rtol=1e-6, atol=3e-7)
    np.testing.assert_allclose(Tg.cpu().numpy(), Rg, rtol=1e-6, atol=3e-7)


def test_sample_poses_rng_distribution(rac):
    import data_loaders
    g = torch.Generator(device="cuda"); g.manual_seed(11)
    Ti, Tg = data_loaders.sample_T_CO_inits_and_gts_device(4096, SCENE_CFG, generator=g)
    g.manual_seed(11)
    Ti2, _ = data_loaders.sample_T_CO_inits_and_gts_device(4096, SCENE_CFG, generator=g)
    assert torch.equal(Ti, Ti2)
    Tg = Tg.cpu().numpy().astype(np.float64)
    Ti = Ti.cpu().numpy().astype(np.float64)
    R = Tg[:, :3, :3]
    np.testing.assert_allclose(R @ R.transpose(0, 2, 1), np.tile(np.eye(3), (4096, 1, 1)), atol=1e-6)
    # object within N(0, 0.1) of the world origin, camera 2.5 +- 0.1 m away: z in camera frame ~ 2.5
    assert abs(Tg[:, 2, 3].mean() - 2.5) < 0.02 and 0.08 < Tg[:, 0, 3].std() < 0.12
    # perturbation: relative rotation angle <= 30 deg, translation offset within the +-0.1 cube
    dR = Ti[:, :3, :3] @ R.transpose(0, 2, 1)
    ang = np.degrees(np.arccos(np.clip((np.trace(dR, axis1=1, axis2=2) - 1) / 2, -1, 1)))
    assert ang.max() <= 30.0 + 1e-3 and ang.mean() > 10.0


@pytest.mark.parametrize("shape", [(48, 64), (33, 47), (240, 320)])
def test_silhouette_matches_oracle(rac, gold, shape):
    rng = np.random.default_rng(shape[0])
    H, W = shape
    B = 3
    imgs = np.where(rng.uniform(size=(B, H, W, 1)) > 0.93, rng.uniform(0.1, 1, (B, H, W, 3)), 0.0).astype(np.float32)
    imgs[0, :2, :3] = 0.5; imgs[1, -1, -1] = 1.0; imgs[2] = 0.0             # corners, and an empty image
    ring = rac.silhouette_device(imgs, True).cpu().numpy()
    paint = rac.silhouette_device(imgs, False).cpu().numpy()
    for b in range(B):
        np.testing.assert_array_equal(ring[b], aux_ref.silhouette(imgs[b], True).astype(np.float32))
        np.testing.assert_array_equal(paint[b], aux_ref.silhouette(imgs[b], False))
    if shape == (48, 64):
        np.testing.assert_array_equal(rac.silhouette_device(gold["img"], True).cpu().numpy(), gold["sil_only"].astype(np.float32))


def test_visualization_module_and_blend(rac, gold):
    import visualization as V
    img = gold["img"].copy()
    assert V.create_silhouette(img, silhouette_only=True).dtype == np.float64
    out = V.create_silhouette(img)                       # paints in place like the reference
    assert out is img
    np.testing.assert_array_equal(img, gold["sil_paint"])
    a, b64 = gold["blend_a"], gold["blend_b64"]
    np.testing.assert_array_equal(V.blend_imgs(a, b64.astype(np.float32), 0.3), gold["blend_f32_03"].astype(np.float32).astype(np.float64))
    got = V.blend_imgs(a, b64, 0.5)
    assert got.dtype == np.float64
    np.testing.assert_allclose(got, gold["blend_mixed_05"], rtol=0, atol=3e-8)   # f32 storage of q/255
    np.testing.assert_allclose(V.blend_imgs(a.astype(np.float64), b64, 0.77), gold["blend_f64_077"], rtol=0, atol=3e-8)
    q = np.round(V.blend_imgs(a.astype(np.float64), b64, 0.77) * 255).astype(np.int64)
    np.testing.assert_array_equal(q, np.round(gold["blend_f64_077"] * 255).astype(np.int64))
    with pytest.raises(AssertionError):
        rac.blend_device(a, b64, 1.5)
