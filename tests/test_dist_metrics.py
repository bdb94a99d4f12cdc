"""compute_angular_dist / compute_transl_dist (ref: loss.py:35-52; called every test iteration at
test_model.py:196-197): the shadow's host versions against golden vectors produced by executing the reference
(tests/golden/make_dist_golden.py), the unclamped-arccos nan case included, and against the reference itself when
/root/reference is importable (build container only)."""
import importlib.util
import os

import numpy as np
import torch

import loss as shadow_loss
from conftest import GOLDEN


def test_against_reference_goldens():
    g = np.load(os.path.join(GOLDEN, "dist_golden.npz"))
    assert np.isnan(g["ang"]).sum() == 1, "the golden set must hold the nan case"
    for Tg, Tp, ang, tr in zip(g["Tg"], g["Tp"], g["ang"], g["tr"]):
        with np.errstate(invalid="ignore"):
            a = shadow_loss.compute_angular_dist(torch.tensor(Tg), torch.tensor(Tp))
        t = shadow_loss.compute_transl_dist(torch.tensor(Tg), torch.tensor(Tp))
        if np.isnan(ang):
            assert np.isnan(a)
        else:
            np.testing.assert_allclose(a, ang, rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(t, tr, rtol=1e-6)
        assert isinstance(t, (np.floating, float)) and np.ndim(a) == 0


def test_only_sample_zero_is_scored_and_inputs_may_carry_grad():
    """The reference reads index 0 of the batch only (loss.py:39-40,46-47) and detaches."""
    Tg = torch.eye(4).repeat(3, 1, 1)
    Tp = torch.eye(4).repeat(3, 1, 1).requires_grad_(True)
    Tp2 = Tp.detach().clone()
    Tp2[1:, :3, 3] = 7.0
    Tp2[1:, :3, :3] = 0.0
    assert shadow_loss.compute_transl_dist(Tg, Tp) == shadow_loss.compute_transl_dist(Tg, Tp2) == 0.0
    assert shadow_loss.compute_angular_dist(Tg, Tp) == shadow_loss.compute_angular_dist(Tg, Tp2) == 0.0


def test_against_reference_live():
    path = "/root/reference/loss.py"
    if not os.path.isfile(path):
        import pytest
        pytest.skip("reference not present on this machine")
    import sys
    sys.path.insert(0, "/root/reference")
    try:
        spec = importlib.util.spec_from_file_location("_ref_loss_live", path)
        ref = importlib.util.module_from_spec(spec)
        saved = sys.modules.pop("rotation_representation", None)      # the reference's loss.py imports its own
        try:
            spec.loader.exec_module(ref)
        finally:
            sys.modules.pop("rotation_representation", None)
            if saved is not None:
                sys.modules["rotation_representation"] = saved
    finally:
        sys.path.remove("/root/reference")
    rng = np.random.default_rng(0)
    for _ in range(20):
        Tg = torch.tensor(rng.normal(size=(2, 4, 4)).astype(np.float32))
        Tp = torch.tensor(rng.normal(size=(2, 4, 4)).astype(np.float32))
        with np.errstate(invalid="ignore"):
            a, b = ref.compute_angular_dist(Tg, Tp), shadow_loss.compute_angular_dist(Tg, Tp)
        assert (np.isnan(a) and np.isnan(b)) or a == b
        assert ref.compute_transl_dist(Tg, Tp) == shadow_loss.compute_transl_dist(Tg, Tp)
