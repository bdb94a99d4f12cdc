"""-m gpu: RenderCompareStep (one refinement iteration replayed as a CUDA graph) against the eager calls of the
shadow modules and against the oracle."""
import numpy as np
import pytest
import torch

from conftest import intrinsics, sample_poses
from oracle import assemble_ref, pose_loss_ref
from oracle import raster_oracle as ro

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rot,loss,C7", [("6D", "add_l1_disentangled", False), ("SVD", "add_l1", True),
                                         ("3D", "add_l1_disentl_scaled", False)])
def test_graph_step_matches_eager_and_oracle(rot, loss, C7):
    import rac
    from rac import mesh as M
    m = M.bumpy_sphere(level=3, seed=4)
    B, W, H, N = 5, 160, 120, 300
    rng = np.random.default_rng(3)
    T, Tg = sample_poses(B, seed=21)
    K = intrinsics(B, W, H)
    D = {"6D": 9, "SVD": 12, "3D": 6}[rot]
    y = (0.1 * rng.normal(size=(B, D))).astype(np.float32)
    if rot == "6D":
        y[:, 0] += 1; y[:, 4] += 1
    elif rot == "SVD":
        y[:, 0] += 1; y[:, 4] += 1; y[:, 8] += 1
    y[:, -1] = 1.0 + 0.05 * rng.normal(size=B)
    pts = m.verts[rng.integers(0, len(m.verts), (B, N))].astype(np.float32)
    gt = rng.uniform(0, 1, (B, H, W, 3)).astype(np.float32)

    st = rac.RenderCompareStep(m, B, (W, H), rot_repr=rot, loss=loss, n_points=N, use_norm_depth=C7)
    st.T_pred.copy_(torch.tensor(T)); st.T_prev.copy_(torch.tensor(T)); st.K.copy_(torch.tensor(K))
    st.gt.copy_(torch.tensor(gt)); st.y.copy_(torch.tensor(y)); st.T_gt.copy_(torch.tensor(Tg)); st.pts.copy_(torch.tensor(pts))
    st.run_eager()
    torch.cuda.synchronize()
    eager = [t.clone() for t in (st.x, st.T_new, st.loss, st.grad_y)]
    for t in (st.x, st.T_new, st.loss, st.grad_y):
        t.zero_()
    st.capture()
    assert st.launches_per_run >= 5            # setup, gt_norm, raster, finalize_live + the fused pose/loss step
    for _ in range(3):                       # replays must be idempotent
        st.run()
    torch.cuda.synchronize()
    assert torch.equal(st.T_new, eager[1]) and torch.equal(st.grad_y, eager[3]) and torch.equal(st.loss, eager[2])
    np.testing.assert_allclose(st.x.cpu().numpy(), eager[0].cpu().numpy(), rtol=1e-6, atol=1e-6)

    # oracle: CPU rasteriser restatement + reference-pinned torch pose/loss
    rgb_o, depth_o, _ = ro.render_batch(m, T, K, (W, H))
    ref = assemble_ref.prepare_model_input(rgb_o, gt.astype(np.float64), depth_o, C7).numpy()
    x = st.x.cpu().numpy()
    c0 = 1 if C7 else 0
    assert (np.abs(x[:, c0:c0 + 3] - ref[:, c0:c0 + 3]).max(1) > 2 / 255).mean() <= 1e-3
    np.testing.assert_allclose(x[:, c0 + 3:], ref[:, c0 + 3:], rtol=1e-5, atol=2e-6)
    yc = torch.tensor(y, requires_grad=True)
    Tn = pose_loss_ref.pose_update(yc, torch.tensor(T), rot, torch.tensor(K))
    if loss == "add_l1":
        lc = pose_loss_ref.add_l1(torch.tensor(Tg), Tn, torch.tensor(pts))
    elif loss == "add_l1_disentangled":
        lc = pose_loss_ref.disentangled(torch.tensor(Tg), Tn, torch.tensor(pts))
    else:
        lc = pose_loss_ref.disentangled_scaled(torch.tensor(T), Tn, torch.tensor(Tg), torch.tensor(pts))
    lc.backward()
    np.testing.assert_allclose(st.loss.item(), lc.item(), rtol=1e-5)
    np.testing.assert_allclose(st.T_new.cpu().numpy(), Tn.detach().numpy(), rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(st.grad_y.cpu().numpy(), yc.grad.numpy(), rtol=1e-3, atol=2e-6)


@pytest.mark.parametrize("rot,loss", [("6D", "add_l1_disentangled"), ("SVD", "add_l1_disentl_scaled")])
def test_serial_chain_three_iterations_against_the_oracle_chain(rot, loss):
    """run_chain(): render -> pose update + loss -> T_pred <- T_new, three times (train_model.py:236-250 with a fixed head
    output).  Iteration-3 poses, loss and the rendered input are checked against the oracle chain."""
    import rac
    from rac import mesh as M
    m = M.bumpy_sphere(level=3, seed=4)
    B, W, H, N = 6, 160, 120, 200
    rng = np.random.default_rng(13)
    T, Tg = sample_poses(B, seed=31)
    K = intrinsics(B, W, H)
    D = {"6D": 9, "SVD": 12}[rot]
    y = (0.05 * rng.normal(size=(B, D))).astype(np.float32)
    y[:, 0] += 1; y[:, 4] += 1
    if rot == "SVD":
        y[:, 8] += 1
    y[:, -1] = 1.0 + 0.02 * rng.normal(size=B)
    pts = m.verts[rng.integers(0, len(m.verts), (B, N))].astype(np.float32)
    gt = rng.uniform(0, 1, (B, H, W, 3)).astype(np.float32)
    st = rac.RenderCompareStep(m, B, (W, H), rot_repr=rot, loss=loss, n_points=N)
    st.T_pred.copy_(torch.tensor(T)); st.T_prev.copy_(torch.tensor(T)); st.K.copy_(torch.tensor(K))
    st.gt.copy_(torch.tensor(gt)); st.y.copy_(torch.tensor(y)); st.T_gt.copy_(torch.tensor(Tg)); st.pts.copy_(torch.tensor(pts))
    st.capture()
    st.T_pred.copy_(torch.tensor(T)); st.T_prev.copy_(torch.tensor(T))       # (capture warm-up does not move the pose; be explicit)
    Tc = torch.tensor(T)
    for it in range(3):
        st.run_chain()
        yc = torch.tensor(y, requires_grad=True)
        Tn = pose_loss_ref.pose_update(yc, Tc, rot, torch.tensor(K))
        if loss == "add_l1_disentangled":
            lc = pose_loss_ref.disentangled(torch.tensor(Tg), Tn, torch.tensor(pts))
        else:
            lc = pose_loss_ref.disentangled_scaled(Tc, Tn, torch.tensor(Tg), torch.tensor(pts))
        lc.backward()
        T_in = Tc.numpy()
        Tc = Tn.detach()
    torch.cuda.synchronize()
    np.testing.assert_allclose(st.T_pred.cpu().numpy(), Tc.numpy(), rtol=1e-5, atol=1e-6)      # iteration-3 poses
    np.testing.assert_allclose(st.loss.item(), lc.item(), rtol=1e-5)
    np.testing.assert_allclose(st.grad_y.cpu().numpy(), yc.grad.numpy(), rtol=1e-3, atol=2e-6)
    # the third render ran under the pose of iteration 2
    rgb_o, depth_o, _ = ro.render_batch(m, T_in, K, (W, H))
    ref = assemble_ref.prepare_model_input(rgb_o, gt.astype(np.float64), depth_o, False).numpy()
    x = st.x.cpu().numpy()
    assert (np.abs(x[:, :3] - ref[:, :3]).max(1) > 2 / 255).mean() <= 2e-3


def test_mixed_mesh_step_is_capturable():
    """A batch that mixes meshes carries a per-image mesh-id table; it reaches the device as kernel parameters, so the
    captured step holds no pointer into host memory and replays bit-identically to the eager run."""
    import rac
    from rac import mesh as M
    ms = [M.bumpy_sphere(level=3, seed=4), M.table(3), M.box(0.7, 0.5, 0.3, 2)]
    B, W, H, N = 7, 160, 120, 100
    batch = [ms[i % 3] for i in range(B)]
    rng = np.random.default_rng(3)
    T, Tg = sample_poses(B, seed=77)
    K = intrinsics(B, W, H)
    y = (0.05 * rng.normal(size=(B, 9))).astype(np.float32)
    y[:, 0] += 1; y[:, 4] += 1; y[:, 8] = 1.0
    st = rac.RenderCompareStep(batch, B, (W, H), rot_repr="6D", loss="add_l1", n_points=N)
    st.T_pred.copy_(torch.tensor(T)); st.K.copy_(torch.tensor(K)); st.gt.copy_(torch.rand((B, H, W, 3)))
    st.y.copy_(torch.tensor(y)); st.T_gt.copy_(torch.tensor(Tg)); st.pts.copy_(torch.tensor(rng.normal(size=(B, N, 3)).astype(np.float32)))
    st.run_eager()
    torch.cuda.synchronize()
    eager = st.x.clone()
    st.x.zero_()
    st.capture()
    for _ in range(2):
        st.run()
    torch.cuda.synchronize()
    assert torch.equal(st.x, eager)
    _, _, rgba_o = ro.render_batch_mixed(batch, T, K, (W, H))
    cov = (st.x[:, :3].abs().sum(1) > 0).cpu().numpy()             # any sample covered (resolved colour > 0)
    assert ((cov != (rgba_o[..., :3].max(-1) > 0)).reshape(B, -1).sum(1) <= 0.002 * W * H).all()


@pytest.mark.parametrize("rot", ["6D", "SVD", "3D"])
@pytest.mark.parametrize("loss", ["add_l1", "add_l2", "add_l1_disentangled", "add_l1_disentl_scaled"])
def test_fused_pose_loss_step_equals_the_reference_named_calls(rot, loss):
    """rac_pose_loss_step (one launch: pose update + loss + batch mean + backward to y) against the four launches behind
    calculate_T_CO_pred / compute_*_loss / backward that the drivers use — same per-sample device code, so the results
    agree to the last bits; f32 and f64 (Q6) input poses, ragged B and N."""
    import loss as L
    import rotation_representation as RR
    from rac.ops import pose_loss_step
    rng = np.random.default_rng(hash((rot, loss)) % 2 ** 31)
    B, N = 37, 333
    D = {"6D": 9, "SVD": 12, "3D": 6}[rot]
    T, Tg = sample_poses(B, seed=3)
    y = (0.2 * rng.normal(size=(B, D))).astype(np.float32)
    if rot != "3D":
        y[:, 0] += 1; y[:, 4] += 1
    if rot == "SVD":
        y[:, 8] += 1
    y[:, -1] = 1.0 + 0.05 * rng.normal(size=B)
    K = torch.tensor(intrinsics(B, 320, 240))
    pts = torch.tensor(rng.normal(size=(B, N, 3)).astype(np.float32) * 0.3, device="cuda")
    T_gt = torch.tensor(Tg, device="cuda")
    for dtype in (torch.float32, torch.float64):
        T_in = torch.tensor(T, device="cuda").to(dtype)
        T_prev = torch.tensor(T, device="cuda")
        yg = torch.tensor(y, device="cuda", requires_grad=True)
        T_new = RR.calculate_T_CO_pred(yg, T_in, rot, K)
        if loss == "add_l1":
            l = L.compute_ADD_L1_loss(T_gt, T_new, pts)
        elif loss == "add_l2":
            import contextlib, io
            with contextlib.redirect_stdout(io.StringIO()):
                l = L.compute_ADD_L2_loss(T_gt, T_new, pts)
        elif loss == "add_l1_disentangled":
            l = L.compute_disentangled_ADD_L1_loss(T_gt, T_new, pts)             # train_model.py:280 order
        else:
            l = L.compute_scaled_disentl_ADD_L1_loss(T_prev, T_new, T_gt, pts)
        l.backward()
        Tn, lf, gy, per = pose_loss_step(torch.tensor(y, device="cuda"), T_in, rot, float(K[0, 0, 0]), float(K[0, 1, 1]), T_gt, pts, loss,
                                         T_prev=T_prev if loss == "add_l1_disentl_scaled" else None)
        torch.cuda.synchronize()
        assert torch.equal(Tn, T_new.detach()), (rot, loss, dtype)
        np.testing.assert_allclose(lf.item(), l.item(), rtol=1e-6)
        np.testing.assert_allclose(gy.cpu().numpy(), yg.grad.cpu().numpy(), rtol=2e-5, atol=1e-8)
        assert per.shape == (B,) and np.isclose(per.mean().item(), l.item(), rtol=1e-5)


def test_step_reference_named_path_still_matches_the_fused_one():
    import rac
    from rac import mesh as M
    m = M.bumpy_sphere(level=3, seed=4)
    B, W, H, N = 5, 160, 120, 200
    rng = np.random.default_rng(2)
    T, Tg = sample_poses(B, seed=8)
    outs = []
    for fused in (True, False):
        st = rac.RenderCompareStep(m, B, (W, H), rot_repr="SVD", loss="add_l1_disentl_scaled", n_points=N, fused_pose_loss=fused)
        r = np.random.default_rng(2)
        y = (0.1 * r.normal(size=(B, 12))).astype(np.float32); y[:, 0] += 1; y[:, 4] += 1; y[:, 8] += 1; y[:, -1] = 1.0
        st.T_pred.copy_(torch.tensor(T)); st.T_prev.copy_(torch.tensor(T)); st.K.copy_(torch.tensor(intrinsics(B, W, H)))
        st.gt.copy_(torch.tensor(r.uniform(0, 1, (B, H, W, 3)).astype(np.float32))); st.y.copy_(torch.tensor(y))
        st.T_gt.copy_(torch.tensor(Tg)); st.pts.copy_(torch.tensor(m.verts[r.integers(0, len(m.verts), (B, N))]))
        st.capture()
        st.T_pred.copy_(torch.tensor(T)); st.T_prev.copy_(torch.tensor(T))
        for _ in range(3):
            st.run_chain()
        torch.cuda.synchronize()
        outs.append([t.clone() for t in (st.x, st.T_pred, st.loss, st.grad_y)])
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][0], outs[1][0])
    np.testing.assert_allclose(outs[0][2].item(), outs[1][2].item(), rtol=1e-6)
    np.testing.assert_allclose(outs[0][3].cpu().numpy(), outs[1][3].cpu().numpy(), rtol=2e-5, atol=1e-8)
