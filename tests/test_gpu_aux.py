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


# ---- 8-bit gt crops (rac_gt_norm_u8): image_dataloaders.py:79-83 + data_loaders.py:79-82 from the raw codes -------
def _u8_crops(B, H, W, seed):
    rng = np.random.default_rng(seed)
    g = rng.integers(0, 256, (B, H, W, 3), dtype=np.uint8)
    g[1] = rng.integers(100, 104, (H, W, 3), dtype=np.uint8)        # narrow range
    g[2] = 0
    g[2, H // 2, W // 2, 1] = 255                                   # one bright value in a black crop
    return g


@pytest.mark.parametrize("use_depth", [False, True])
@pytest.mark.parametrize("W,H", [(320, 240), (99, 65), (32, 8)])
def test_gt_u8_crop_matches_reference_cast(rac, use_depth, W, H):
    """uint8 crop in, planes out == (normalize_img(u8 / 255.0)).astype(f32) of the oracle, bit for bit (exact integer
    statistics + a per-image table evaluated in f64); the render / depth planes equal the f32-crop path's."""
    from oracle import assemble_ref
    B = 5
    g = _u8_crops(B, H, W, seed=W)
    rng = np.random.default_rng(1)
    init = rng.uniform(0, 1, (B, H, W, 3)).astype(np.float32)
    depth = np.where(rng.uniform(size=(B, H, W)) < 0.3, rng.uniform(2.0, 3.0, (B, H, W)), 0.0).astype(np.float32)
    ref = assemble_ref.prepare_model_input(init, g.astype(np.float64) / 255.0, depth, use_depth).numpy()
    c0 = 1 if use_depth else 0
    for form in ("numpy", "device", "pinned"):
        gt = g if form == "numpy" else torch.from_numpy(g).cuda() if form == "device" else torch.from_numpy(g).pin_memory()
        x = rac.assemble_input_device(init, gt, depth if use_depth else None, use_depth).cpu().numpy()
        assert x.shape == ref.shape
        a, b = x[:, c0 + 3:], ref[:, c0 + 3:]
        same = a.view(np.uint32) == b.view(np.uint32)
        assert same.mean() > 0.9999, (form, same.mean())            # a table entry on a rounding boundary: <= 1 ulp
        np.testing.assert_allclose(a, b, rtol=2e-7, atol=0)
        x32 = rac.assemble_input_device(init, (g.astype(np.float64) / 255.0).astype(np.float32), depth if use_depth else None,
                                        use_depth).cpu().numpy()
        assert np.array_equal(x[:, :c0 + 3], x32[:, :c0 + 3])
        # (full-range crops only: an f32 crop of a 4-level image has lost more than 1e-5 of its spread before any kernel runs)
        np.testing.assert_allclose(a[[0, 3, 4]], x32[[0, 3, 4], c0 + 3:], rtol=1e-5, atol=2e-6)


def test_gt_u8_constant_crop_is_nan_like_the_reference(rac):
    """normalize_img has no zero-std guard (data_loaders.py:79-82): a constant crop is 0/0 = nan everywhere."""
    g = np.full((2, 16, 32, 3), 77, dtype=np.uint8)
    g[1, 0, 0, 0] = 78
    init = np.zeros((2, 16, 32, 3), dtype=np.float32)
    x = rac.assemble_input_device(init, g, None, False).cpu().numpy()
    assert np.isnan(x[0, 3:]).all() and np.isfinite(x[1, 3:]).all() and np.array_equal(x[:, :3], np.zeros_like(x[:, :3]))


def test_gt_u8_full_size_through_the_shadow_module(rac):
    """data_loaders.prepare_model_input with a pinned uint8 crop at the bench size (B=64, 320x240): statistics of every
    image are those of a normalised image (mean 0, std 1), planes equal the f32-crop path within the f32 bar, and the
    call is repeatable (staging ring reuse)."""
    import data_loaders as DL
    B, H, W = 64, 240, 320
    g = torch.from_numpy(np.random.default_rng(4).integers(0, 256, (B, H, W, 3), dtype=np.uint8)).pin_memory()
    init = torch.rand((B, H, W, 3), device="cuda")
    from rac import render as R
    prev = R.set_gt_cache(False)        # every call uploads and normalises
    try:
        x1 = DL.prepare_model_input(init, g, None, False)
        x2 = DL.prepare_model_input(init, g, None, False)
    finally:
        R.set_gt_cache(prev)
    x3 = DL.prepare_model_input(init, (g.double() / 255.0).float().pin_memory(), None, False)
    assert torch.equal(x1, x2)
    gt = x1[:, 3:].double()
    assert gt.mean(dim=(1, 2, 3)).abs().max() < 1e-6 and (gt.std(dim=(1, 2, 3), unbiased=False) - 1).abs().max() < 1e-6
    assert torch.equal(x1[:, :3], x3[:, :3])
    torch.testing.assert_close(x1[:, 3:], x3[:, 3:], rtol=1e-5, atol=2e-6)


def test_first_sample_intrinsics_under_nccl_with_cpu_cam_mats(monkeypatch):
    """The reference's cam_mats live on the CPU (Q5) and INTEGRATION.md tells users to pass exactly that; NCCL has no CPU
    backend, so the broadcast buffer must be staged on this rank's GPU.  One-rank NCCL group, broadcast path forced."""
    import torch.distributed as dist
    from rac import dist as rd
    if dist.is_initialized():
        pytest.skip("a process group already exists in this process")
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29541", rank=0, world_size=1,
                            device_id=torch.device("cuda", torch.cuda.current_device()))
    try:
        assert rd._collective_device().type == "cuda"
        monkeypatch.setattr(rd, "world", lambda: (0, 2))           # take the ws > 1 branch: a real NCCL broadcast
        Ks = torch.tensor([[[444.0, 0, 160], [0, 333.0, 120], [0, 0, 1]]], dtype=torch.float64).repeat(3, 1, 1)      # CPU, f64
        assert rd.first_sample_intrinsics(Ks) == (444.0, 333.0)
        assert rd.first_sample_intrinsics(Ks.cuda()) == (444.0, 333.0)
    finally:
        dist.destroy_process_group()


def test_bulk_copy_gt_kernel_is_bit_identical_to_the_register_staged_one(tmp_path):
    """gt_norm_tma_kernel (cp.async.bulk through a shared-memory ring; the standalone default) against gt_norm_kernel
    (RAC_GT_TMA=0, read once per process: a child process) on ragged shapes: slices shorter than one 1024-px chunk, sizes
    that are not a multiple of the chunk, B = 1, and the bench shape."""
    import hashlib
    import subprocess
    import sys
    script = tmp_path / "gt_variants.py"
    script.write_text('''
import hashlib, sys
sys.path[:0] = %r
import numpy as np, torch, rac
from rac import render as R
R.set_gt_cache(False)
h = hashlib.sha1()
for (B, H, W, use_depth) in [(1, 8, 12, False), (3, 40, 52, True), (5, 120, 160, False), (2, 240, 320, True), (9, 67, 100, False)]:
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + W)
    init = torch.rand((B, H, W, 3), device="cuda", generator=g)
    gt = torch.rand((B, H, W, 3), device="cuda", generator=g)
    depth = torch.rand((B, H, W), device="cuda", generator=g) * 3
    for rep in range(2):
        x = rac.assemble_input_device(init, gt, depth if use_depth else None, use_depth)
    h.update(x.cpu().numpy().tobytes())
print(h.hexdigest())
''' % (list(sys.path[:4]),))
    import os
    outs = {}
    for tma in ("1", "0"):
        r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, env=dict(os.environ, RAC_GT_TMA=tma))
        assert r.returncode == 0, r.stderr[-2000:]
        outs[tma] = r.stdout.strip().splitlines()[-1]
    assert outs["1"] == outs["0"], outs
