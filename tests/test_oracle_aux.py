"""Pins oracle/aux_ref.py against golden vectors made with the libraries the reference calls
(scipy.signal.convolve2d, PIL.Image.blend) and with this repo's se3_helpers shadow under a recorded RNG
(tests/golden/make_aux_golden.py), and — where those libraries import — against them directly."""
import os

import numpy as np
import pytest

from conftest import SCENE_CFG
from oracle import aux_ref

GOLD = os.path.join(os.path.dirname(__file__), "golden", "aux_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return {k: v for k, v in np.load(GOLD).items()}


def test_silhouette_matches_golden(gold):
    np.testing.assert_array_equal(aux_ref.silhouette(gold["img"], True), gold["sil_only"])
    np.testing.assert_array_equal(aux_ref.silhouette(gold["img"], False), gold["sil_paint"])
    assert gold["sil_only"].sum() > 100 and set(np.unique(gold["sil_only"])) == {0.0, 1.0}


def test_blend_matches_golden(gold):
    a, b64 = gold["blend_a"], gold["blend_b64"]
    np.testing.assert_array_equal(aux_ref.blend(a, b64.astype(np.float32), 0.3), gold["blend_f32_03"])
    np.testing.assert_array_equal(aux_ref.blend(a, b64, 0.5), gold["blend_mixed_05"])
    np.testing.assert_array_equal(aux_ref.blend(a.astype(np.float64), b64, 0.77), gold["blend_f64_077"])


def test_poses_match_golden_and_are_rigid(gold):
    Ti, Tg = aux_ref.poses_from_variates(gold["u01"], gold["n01"], SCENE_CFG)
    np.testing.assert_allclose(Ti, gold["T_init"], rtol=0, atol=2e-7)
    np.testing.assert_allclose(Tg, gold["T_gt"], rtol=0, atol=2e-7)
    for T in (Ti, Tg):
        R = T[:, :3, :3].astype(np.float64)
        np.testing.assert_allclose(R @ R.transpose(0, 2, 1), np.tile(np.eye(3), (len(T), 1, 1)), atol=1e-6)
        np.testing.assert_allclose(np.linalg.det(R), 1.0, atol=1e-6)
        assert np.all(np.abs(T[:, 2, 3] - 2.5) < 0.1 + 0.6)          # camera 2.5 +- 0.1 m in front, object near the origin
        np.testing.assert_array_equal(T[:, 3], np.tile([0, 0, 0, 1], (len(T), 1)))


def test_against_the_reference_libraries_directly():
    sig = pytest.importorskip("scipy.signal")
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(9)
    img = np.where(rng.uniform(size=(33, 47, 1)) > 0.9, rng.uniform(0.1, 1, (33, 47, 3)), 0.0)
    sil = np.where(np.mean(img, 2) > 0, 1.0, 0.0)
    ring = np.clip(sig.convolve2d(sil, np.ones((5, 5)), mode="same"), 0, 1) - sil
    np.testing.assert_array_equal(aux_ref.silhouette(img, True), ring)
    a, b = rng.uniform(0, 1, (33, 47, 3)), rng.uniform(0, 1, (33, 47, 3)).astype(np.float32)
    for alpha in (0.0, 0.25, 0.5, 0.9, 1.0):
        ref = np.asarray(Image.blend(Image.fromarray(np.uint8(a * 255.0)), Image.fromarray(np.uint8(b * 255.0)), alpha)) / 255.0
        np.testing.assert_array_equal(aux_ref.blend(a, b, alpha), ref)
