"""-m gpu: multi-part (.glb) meshes through the reference API (renderer.render_scene / data_loaders.render_batch) against
the CPU oracle fed the same primitives, and the bounded device mesh cache."""
import numpy as np
import pytest
import torch

from conftest import intrinsics, sample_poses
from oracle import raster_oracle as ro
from test_gpu_raster import compare

pytestmark = pytest.mark.gpu


def _scene_prims():
    from rac import mesh as M
    a = M.with_texture(M.bumpy_sphere(level=3, seed=5), size=128, seed=2)
    b = M.box(0.5, 0.3, 0.2, 2)
    c = M.icosphere(2, radius=0.2)
    rng = np.random.default_rng(1)
    vc = np.concatenate([rng.integers(30, 255, (len(c.verts), 3)), np.full((len(c.verts), 1), 255)], 1).astype(np.uint8)
    return [dict(verts=a.verts, faces=a.faces, uv=a.uv, tex=a.tex, material=(1, 1, 1, 1, 0.1, 0.7)),
            dict(verts=b.verts + np.float32([0.45, 0.1, 0.0]), faces=b.faces, material=(0.8, 0.2, 0.1, 1, 0.5, 0.4)),
            dict(verts=c.verts + np.float32([-0.45, 0.0, 0.2]), faces=c.faces, vcolor=vc)]


def test_glb_scene_matches_oracle(tmp_path):
    """A 3-primitive GLB (textured, factor-coloured, COLOR_0 without a material) in one depth buffer: coverage / depth
    bit-level bars and colour within 2/255, through render_batch_device and through the reference-named entry points."""
    import rac
    import data_loaders as DL
    import renderer
    from rac import gltf
    from rac import mesh as M
    path = str(tmp_path / "scene.glb")
    gltf.save_glb(path, _scene_prims(), node_translation=(3, 0, 0))
    m = M.load_mesh(path)
    B, W, H = 6, 320, 240
    T, _ = sample_poses(B, seed=12)
    K = intrinsics(B, W, H)
    out = rac.render_batch_device(T, path, K, (W, H), want_rgba8=True)
    orc = ro.render_batch(m, T, K, (W, H))
    compare(out, orc, W, H, "glb")
    q = out["rgba8"].cpu().numpy().astype(int)
    cov = q[..., :3][out["depth"].cpu().numpy() > 0]
    assert ((cov[:, 0] > cov[:, 1] + 50) & (cov[:, 0] > cov[:, 2] + 50)).sum() > 200, "the factor-coloured box must show"
    # the reference-named entry points (square images: renderer.py:78)
    S = 200
    Ks = intrinsics(1, S, S)
    img, depth = renderer.render_scene(path, T[0], K=Ks[0], img_size=S)
    rgb_o, depth_o, _ = ro.render_batch(m, T[:1], Ks, S)
    assert img.shape == (S, S, 3) and ((depth > 0) != (depth_o[0] > 0)).mean() <= 1e-3
    assert (np.abs(img - rgb_o[0]).max(-1) > 2 / 255).mean() <= 1e-3
    imgs, depths = DL.render_batch(T[:2], [path, path], np.repeat(Ks, 2, 0), S)
    assert np.asarray(imgs).shape == (2, S, S, 3)
    np.testing.assert_array_equal(np.asarray(depths)[0], depth)
    # the vertex source of the loss points (data_loaders.py:19-32) sees the rotated, merged vertices
    assert DL.get_vertices(path).shape == (len(m.verts), 3)


def test_glb_in_a_mixed_batch_and_fused_concat(tmp_path):
    import rac
    from rac import gltf
    from rac import mesh as M
    from oracle import assemble_ref
    path = str(tmp_path / "scene.glb")
    gltf.save_glb(path, _scene_prims())
    glb = M.load_mesh(path)
    plain = M.bumpy_sphere(level=3, seed=9)
    batch = [glb, plain, glb, plain, glb]
    B, W, H = len(batch), 160, 120
    T, _ = sample_poses(B, seed=4)
    K = intrinsics(B, W, H)
    gt = np.random.default_rng(0).uniform(0, 1, (B, H, W, 3)).astype(np.float32)
    x, depth, rgb = rac.render_concat_device(T, batch, K, (W, H), gt, use_norm_depth=True, want_depth=True, want_rgb=True)
    rgb_o, depth_o, _ = ro.render_batch_mixed(batch, T, K, (W, H))
    ref = assemble_ref.prepare_model_input(rgb_o, gt.astype(np.float64), depth_o, True).numpy()
    x = x.cpu().numpy()
    assert ((depth.cpu().numpy() > 0) != (depth_o > 0)).mean() <= 1e-3
    assert (np.abs(x[:, 1:4] - ref[:, 1:4]).max(1) > 2 / 255).mean() <= 1e-3
    np.testing.assert_allclose(x[:, 4:], ref[:, 4:], rtol=1e-5, atol=2e-6)


def test_device_mesh_cache_is_bounded(monkeypatch):
    """More distinct meshes than the device table holds (kMaxMeshes = 4096) cycle through render_batch: the LRU cache evicts
    (rac_mesh_destroy) instead of dying with 'mesh table full', and an evicted mesh renders correctly when it comes back."""
    import rac
    from rac import mesh as M
    M.clear_cache()
    rng = np.random.default_rng(0)
    n_meshes, B, S = 4200, 70, 64
    base = M.box(0.5, 0.4, 0.3, 1)
    meshes = [M.Mesh(base.verts * np.float32(0.6 + 0.8 * rng.uniform()), base.faces) for _ in range(n_meshes)]
    T, _ = sample_poses(B, seed=2)
    K = intrinsics(B, S, S)
    first = None
    for i in range(0, n_meshes, B):
        out = rac.render_batch_device(T, meshes[i:i + B], K, S, want_rgb=False)
        if first is None:
            first = out["depth"].clone()
    assert len(M._cache) <= M.MAX_DEVICE_MESHES
    again = rac.render_batch_device(T, meshes[:B], K, S, want_rgb=False)["depth"]      # long evicted: uploaded again
    assert torch.equal(first, again)
    M.clear_cache()


def test_device_mesh_held_by_a_caller_is_never_evicted():
    """A DeviceMesh somebody still references (rac.RenderCompareStep.meshes, the owner of a captured graph: mesh ids are
    baked into graphs by value and a freed table slot is reused by the next upload) survives any amount of cache
    churn with its handle intact; unreferenced entries around it are evicted in batches."""
    import ctypes
    import rac
    from rac import _lib, mesh as M
    M.clear_cache()
    rng = np.random.default_rng(1)
    B, S = 64, 64
    base = M.box(0.5, 0.4, 0.3, 1)
    mk = lambda: M.Mesh(base.verts * np.float32(0.6 + 0.8 * rng.uniform()), base.faces)
    T, _ = sample_poses(B, seed=4)
    K = intrinsics(B, S, S)
    held_host = [mk() for _ in range(B)]
    held = [M.device_mesh(m) for m in held_host]                 # what a loop-owning caller keeps
    handles = [d.handle for d in held]
    first = rac.render_batch_device(T, held, K, S, want_rgb=False)["depth"].clone()
    for _ in range(0, M.MAX_DEVICE_MESHES + 4 * B, B):           # churn well past the bound
        rac.render_batch_device(T, [mk() for _ in range(B)], K, S, want_rgb=False)
    assert len(M._cache) <= M.MAX_DEVICE_MESHES
    v, t = ctypes.c_int(), ctypes.c_int()
    for d, h in zip(held, handles):
        assert d.handle == h and M._cache.get(("obj", id(d.mesh))) is d
        assert _lib.lib().rac_mesh_info(h, ctypes.byref(v), ctypes.byref(t)) == 0 and (v.value, t.value) == (d.V, d.T)
    assert torch.equal(first, rac.render_batch_device(T, held, K, S, want_rgb=False)["depth"])
    del held, d
    M.clear_cache()


def test_double_sided_open_mesh_matches_oracle(tmp_path):
    """An OPEN textured surface (a bumpy sphere with 40 % of its faces removed) whose material is doubleSided: through the
    holes the inside of the far wall is visible — drawn back faces, unlit, with the right texture coordinates (the
    attribute order of an unswapped record).  Coverage / depth / colour bars against the oracle; and the same file
    without doubleSided must show fewer covered pixels."""
    import rac
    from rac import gltf
    from rac import mesh as M
    body = M.with_texture(M.bumpy_sphere(level=4, seed=3, decimate_to=3000), size=64, seed=1)
    B, W, H = 4, 240, 200
    T, _ = sample_poses(B, seed=6)
    K = intrinsics(B, W, H)
    cov = {}
    for ds in (True, False):
        path = str(tmp_path / ("open_%d.glb" % ds))
        gltf.save_glb(path, [dict(verts=body.verts, faces=body.faces, uv=body.uv, tex=body.tex, material=(1, 1, 1, 1, 0.1, 0.6), double_sided=ds)])
        m = M.load_mesh(path)
        out = rac.render_batch_device(T, path, K, (W, H), want_rgba8=True)
        compare(out, ro.render_batch(m, T, K, (W, H)), W, H, "double_sided=%s" % ds)
        cov[ds] = (out["depth"] > 0).float().mean().item()
    assert cov[True] > cov[False] * 1.1, cov
