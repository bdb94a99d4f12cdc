"""CPU (-m "not gpu"): the raster oracle (oracle/raster_ref.c) against its committed golden vectors, analytic
known answers (SURVEY.md §8c self-consistency checks) and the numpy restatement of prepare_model_input.
Parity of step 1 is UNPINNED with respect to real pyrender (see oracle/raster_ref.c); these tests pin the
oracle to the written spec (R1-R9) and to itself."""
import os

import numpy as np
import pytest

from conftest import intrinsics, sample_poses
from oracle import assemble_ref
from oracle import raster_oracle as ro

GOLD = os.path.join(os.path.dirname(__file__), "golden", "raster_golden.npz")


def _scenes():
    import make_raster_golden as G
    return G.scenes(), G.W, G.H


def test_oracle_reproduces_golden_vectors():
    gold = np.load(GOLD)
    scenes, W, H = _scenes()
    for name, (m, T, K) in scenes.items():
        np.testing.assert_array_equal(gold[f"{name}_T"], T)
        rgb, depth, rgba8 = ro.render_batch(m, T, K, (W, H))
        # libm differences between hosts may move a colour by 1 LSB on a handful of pixels; depth/coverage are exact
        np.testing.assert_array_equal(depth, gold[f"{name}_depth"])
        dq = np.abs(rgba8.astype(int) - gold[f"{name}_rgba8"].astype(int))
        assert dq.max() <= 1 and (dq > 0).mean() < 1e-3, name


def _quad(size=1.0):
    from rac import mesh as M
    v = np.array([[-size, -size, 0], [size, -size, 0], [size, size, 0], [-size, size, 0]], np.float32)
    # object normal -z faces a camera at the origin when the object sits at +z (OpenCV frame)
    f = np.array([[0, 2, 1], [0, 3, 2]], np.int32)
    return M.Mesh(v, f)


def _r7_colour(p, n, base=0.3, metallic=0.2, rough=0.8, ambient=0.1, intensity=15.0):
    """SURVEY.md §8a R7 in float64 for a head-light at the origin; p camera-frame point, n unit normal."""
    dist = np.linalg.norm(p)
    l = -p / dist
    ndl = float(n @ l)
    nl = min(max(ndl, 0.001), 1.0); nv = min(max(abs(ndl), 0.001), 1.0); nh = min(max(ndl, 0.0), 1.0)
    f0 = 0.04 + (base - 0.04) * metallic
    cdiff = base * 0.96 * (1 - metallic)
    a = max(min(rough, 1.0), 0.04) ** 2
    a2 = a * a
    D = a2 / (np.pi * (nh * nh * (a2 - 1) + 1) ** 2)
    k = (rough + 1) ** 2 / 8
    G = (nl / (nl * (1 - k) + k)) * (nv / (nv * (1 - k) + k))
    spec = f0 * G * D / (4 * nl * nv + 1e-4)
    diff = (1 - f0) * cdiff / np.pi
    las = 1.0 / max(0.001, np.cos(np.pi / 9) - np.cos(np.pi / 2))
    lao = -np.cos(np.pi / 2) * las
    cd = min(max(p[2] / dist, 0.0), 1.0)
    att = min(max(cd * las + lao, 0.0), 1.0) ** 2 / dist ** 2
    col = nl * att * intensity * (diff + spec) + base * ambient
    return int(min(max(col, 0.0) ** (1 / 2.2), 1.0) * 255 + 0.5)


@pytest.mark.parametrize("d", [1.5, 2.5, 4.0])
def test_fronto_parallel_quad_known_answer(d):
    """Constant depth d over the quad, closed-form R7 colour, coverage = the projected square."""
    W, H = 160, 120
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, 2, 3] = d
    K = intrinsics(1, W, H)
    size = 0.2
    rgb, depth, rgba8 = ro.render_batch(_quad(size), T, K, (W, H))
    fx, cx, cy = K[0, 0, 0], K[0, 0, 2], K[0, 1, 2]
    half = fx * size / d
    us, vs = np.arange(W) + 0.5, np.arange(H) + 0.5
    inside = (np.abs(vs[:, None] - cy) < half - 1) & (np.abs(us[None, :] - cx) < half - 1)
    outside = (np.abs(vs[:, None] - cy) > half + 1) | (np.abs(us[None, :] - cx) > half + 1)
    assert inside.sum() > 100
    assert (depth[0][inside] > 0).all() and (depth[0][outside] == 0).all()
    np.testing.assert_allclose(depth[0][inside], d, rtol=1e-5)          # 24-bit depth buffer + f32 read-back (R6)
    n = np.array([0.0, 0.0, -1.0])
    for (r, c) in [(H // 2, W // 2), (H // 2 - 5, W // 2 + 7), (H // 2 + 9, W // 2 - 11)]:
        if not inside[r, c]:
            continue
        p = np.array([(c + 0.5 - cx) / fx * d, (r + 0.5 - cy) / fx * d, d])
        q = _r7_colour(p, n)
        assert abs(int(rgba8[0, r, c, 0]) - q) <= 1 and rgba8[0, r, c, 0] == rgba8[0, r, c, 1] == rgba8[0, r, c, 2]
        assert rgba8[0, r, c, 3] == 255
    # R9: max-normalised f32 output
    assert rgb.max() == pytest.approx(1.0)
    np.testing.assert_allclose(rgb[0], rgba8[0, ..., :3].astype(np.float32) / np.float32(rgba8[0, ..., :3].max()), rtol=1e-6)


def test_back_faces_are_culled():
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, 2, 3] = 2.0
    T[0, :3, :3] = np.diag([1.0, -1.0, -1.0])      # flip the quad: its normal now points away from the camera
    _, depth, _ = ro.render_batch(_quad(0.2), T, intrinsics(1, 96, 72), (96, 72))
    assert (depth == 0).all()


def test_sphere_depth_profile_and_silhouette():
    """Icosphere radius r at distance d: silhouette radius fx*r/sqrt(d^2-r^2); centre depth d - r."""
    from rac import mesh as M
    W = H = 200
    r, d = 0.5, 2.5
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, 2, 3] = d
    K = intrinsics(1, W, H)
    _, depth, _ = ro.render_batch(M.icosphere(5, r), T, K, (W, H))
    fx = K[0, 0, 0]
    assert depth[0, H // 2, W // 2] == pytest.approx(d - r, rel=2e-3)
    rad = fx * r / np.sqrt(d * d - r * r)
    area = (depth[0] > 0).sum()
    assert area == pytest.approx(np.pi * rad * rad, rel=0.02)
    # un-projection of covered pixels lands on the sphere
    v, u = np.nonzero(depth[0] > 0)
    z = depth[0][v, u]
    P = np.stack([(u + 0.5 - W / 2) / fx * z, (v + 0.5 - H / 2) / fx * z, z - d], 1)
    rr = np.linalg.norm(P, axis=1)
    assert np.percentile(np.abs(rr - r), 95) < 0.01


def test_msaa_edges_are_fractional_and_interior_is_flat_per_triangle():
    T, _ = sample_poses(1, seed=3)
    from rac import mesh as M
    _, depth, rgba8 = ro.render_batch(M.box(0.6, 0.4, 0.3, 1), T, intrinsics(1, 128, 96), (128, 96))
    a = rgba8[0, ..., 3]
    assert set(np.unique(a)).issubset({0, 64, 128, 191, 255}) and (a == 255).sum() > 100
    assert ((a > 0) & (a < 255)).sum() > 10          # silhouette pixels carry partial coverage (R5)
    assert ((depth[0] > 0) <= (a > 0)).all()          # depth comes from sample 0 only


@pytest.mark.parametrize("use_depth", [False, True])
def test_prepare_model_input_c_matches_numpy_restatement(use_depth):
    rng = np.random.default_rng(0)
    B, H, W = 3, 24, 32
    init = rng.uniform(0, 1, (B, H, W, 3)).astype(np.float32)
    gt = rng.uniform(0, 1, (B, H, W, 3)).astype(np.float32)
    depth = np.where(rng.uniform(size=(B, H, W)) > 0.5, rng.uniform(2, 3, (B, H, W)), 0).astype(np.float32)
    depth[1] = 0.0                                     # Q10: max <= 0.1 -> returned unchanged
    x = ro.prepare_model_input_c(init, gt, depth, use_depth)
    ref = assemble_ref.prepare_model_input(init, gt.astype(np.float64), depth, use_depth).numpy()
    np.testing.assert_allclose(x, ref, rtol=1e-5, atol=1e-6)
    assert x.shape == (B, 7 if use_depth else 6, H, W)


# ---- second source for the (?) items of R5 / R7 / R8: closed-form known answers (tests/known_answers.py) -------------
def _oracle_render(mesh, T, K, wh):
    _, depth, rgba8 = ro.render_batch(mesh, T, K, wh)
    return depth, rgba8


def test_brdf_closed_form_points():
    import known_answers as KA
    KA.check_brdf_points(_oracle_render)
    # the glTF specification's Smith-GGX visibility instead of the shader's Schlick k would move colours by at most this
    assert 0 < KA.spec_visibility_difference() <= 40


def test_resolve_is_the_mean_of_four_samples_and_depth_is_sample_zero():
    import known_answers as KA
    KA.check_half_covered_pixels(_oracle_render)


def test_texture_lod_of_a_minified_checker():
    import known_answers as KA
    KA.check_minified_checker(_oracle_render)


def test_near_plane_clips_per_sample():
    import known_answers as KA
    KA.check_near_plane_clips_per_sample(_oracle_render)
