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
    assert st.launches_per_run >= 7
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
