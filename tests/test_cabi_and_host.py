"""CPU (-m "not gpu"): the C-ABI library loads and exports every symbol include/rac.h declares (no compute call
is made without a GPU), plus the host-side logic of the shadow modules (mesh I/O, pose sampling, LazyArray,
intrinsics) and the no-fallback rule."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rac.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"RAC_API\s+[\w\s\*]+?\b(rac_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from rac import _lib
    names = _declared()
    assert len(names) >= 20, names
    assert sorted(_lib.SIGNATURES) == names, "rac/_lib.py SIGNATURES and include/rac.h disagree"
    L = _lib.lib()                       # dlopen + bind; builds in-tree if the .so is absent
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert getattr(raw, n) is not None
    assert L.rac_version() >= 100
    assert L.rac_launch_count() >= 0


def test_defaults_and_workspace_sizing_without_a_gpu():
    from rac import _lib
    L = _lib.lib()
    s, m = _lib.RacScene(), _lib.RacMaterial()
    L.rac_scene_defaults(ctypes.byref(s))
    L.rac_material_defaults(ctypes.byref(m))
    assert (s.ambient, s.light_intensity, s.znear, s.zfar) == pytest.approx((0.1, 15.0, 0.05, 100.0))   # renderer.py:54,97
    assert s.inner_cone == pytest.approx(np.pi / 9) and s.outer_cone == pytest.approx(np.pi / 2)        # renderer.py:55-56
    assert list(m.base_color) == pytest.approx([0.3, 0.3, 0.3, 1.0]) and m.metallic == pytest.approx(0.2)
    assert L.rac_render_workspace_bytes(0, 320, 240, 100) == 0
    a = L.rac_render_workspace_bytes(8, 320, 240, 10000)
    b = L.rac_render_workspace_bytes(64, 320, 240, 10000)
    assert 0 < a < b and b % 256 == 0
    assert b >= 64 * (10000 * 48 + 320 * 240 * 8)       # records + RGBA8/depth planes at least
    assert L.rac_assemble_workspace_bytes(64) >= 64 * 64


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    """Without a CUDA device every compute entry point fails loudly; nothing is routed to the oracle."""
    import rac
    from rac import _lib, mesh as M
    with pytest.raises(_lib.RacError):
        rac.render_batch_device(np.eye(4, dtype=np.float32)[None], M.icosphere(1), np.eye(3, dtype=np.float32)[None], 32)
    import rotation_representation as RR
    with pytest.raises(_lib.RacError):
        RR.compute_rotation_matrix_from_ortho6d(torch.randn(2, 6))
    src = ""
    pkg = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(".py"):
                src += open(os.path.join(dp, f)).read()
    assert "oracle" not in src.replace("# oracle", ""), "the product package must not import the oracle"


def test_ply_roundtrip_and_procedural_meshes(tmp_path):
    from rac import mesh as M
    m = M.bumpy_sphere(level=2, seed=1)
    rng = np.random.default_rng(0)
    m.vcolor = np.concatenate([rng.integers(0, 255, (len(m.verts), 3)), np.full((len(m.verts), 1), 255)], 1).astype(np.uint8)
    for binary in (True, False):
        p = str(tmp_path / ("m%d.ply" % binary))
        M.save_ply(p, m, binary=binary)
        r = M.load_mesh(p)
        np.testing.assert_allclose(r.verts, m.verts, rtol=1e-6, atol=1e-7)
        np.testing.assert_array_equal(r.faces, m.faces)
        np.testing.assert_array_equal(r.vcolor, m.vcolor)
    assert len(M.icosphere(3).faces) == 20 * 4 ** 3
    assert len(M.bumpy_sphere(level=5, seed=1, decimate_to=10000).faces) == 10000
    t = M.with_texture(M.icosphere(2), size=32)
    assert t.uv.shape == (len(t.verts), 2) and t.tex.shape == (32, 32, 3)
    with pytest.raises(ValueError):
        M.load_mesh(str(tmp_path / "x.obj"))


def test_pose_sampling_follows_se3_helpers():
    """ref: se3_helpers.py:39-88 — gt at distance 2.5 +- 0.1 looking at the origin, init = small perturbation."""
    import se3_helpers as S
    cfg = dict(distance_cam_to_world=2.5, distance_cam_to_world_deviation=0.1, world_to_object_gt_transl_deviation=0.1,
               world_to_object_transl_deviation=0.1, world_to_object_angle_deviation=30)
    np.random.seed(0)
    for _ in range(20):
        Ti, Tg = S.get_T_CO_init_and_gt(cfg)
        A, Bm = Ti.data[0], Tg.data[0]
        for T in (A, Bm):
            np.testing.assert_allclose(T[:3, :3] @ T[:3, :3].T, np.eye(3), atol=1e-9)
            assert np.linalg.det(T[:3, :3]) == pytest.approx(1.0)
            assert 1.8 < T[2, 3] < 3.2
        D = A @ np.linalg.inv(Bm)
        ang = np.degrees(np.arccos(np.clip((np.trace(D[:3, :3]) - 1) / 2, -1, 1)))
        assert ang <= 30.0 + 1e-6
    T_CW = S.get_T_CW(2.5, 0.0)
    np.testing.assert_allclose(np.linalg.norm(np.linalg.inv(T_CW.data[0])[:3, 3]), 2.5, rtol=1e-9)


def test_camera_matrix_and_lazy_array():
    import renderer
    K = renderer.get_camera_matrix(dict(focal_length=50, sensor_width=36, image_resolution=320))   # renderer.py:17-25
    np.testing.assert_allclose(K, [[444.4444444, 0, 160], [0, 444.4444444, 160], [0, 0, 1]], rtol=1e-9)
    from rac.lazy import LazyArray
    t = torch.arange(24, dtype=torch.float32).reshape(2, 3, 4)
    a = LazyArray(t)
    assert a.shape == (2, 3, 4) and len(a) == 2 and a.dtype == np.float32 and a.ndim == 3
    np.testing.assert_array_equal(np.asarray(a), t.numpy())
    np.testing.assert_array_equal(a[1], t.numpy()[1])
    assert float(np.max(a)) == 23.0 and (a / 2)[0, 0, 2] == 1.0
    assert torch.tensor(np.asarray(a)).shape == (2, 3, 4)      # train_model.py:241 pattern


def test_bench_reference_arm_prints_one_json_line():
    """bench.py --impl reference (the oracle port on the host cores): stdout is exactly one JSON line with the
    contract's keys; under a multi-rank launch only rank 0 runs and the other ranks exit 0 without output."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    res = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, res.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "renders/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["steps"] == 1 and d["n_gpus"] == 1 and "workload" in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    env2 = dict(env, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    res2 = subprocess.run(cmd, capture_output=True, text=True, env=env2, timeout=600)
    assert res2.returncode == 0 and res2.stdout.strip() == "", (res2.returncode, res2.stdout, res2.stderr[-500:])


def test_host_pack_recovers_png_codes_exactly():
    """rac_host_pack_gt_u8 (host-only staging, csrc/host_pack.cu): real.png / 255.0 (image_dataloaders.py:79-83) -> the
    PNG's codes, accepted only when EVERY value is exactly q / 255.0; ragged sizes, both source types, thread counts."""
    import ctypes
    from rac import _lib
    L = _lib.lib()
    rng = np.random.default_rng(0)
    ok = ctypes.c_int(-1)
    for n in (1, 63, 64, 65, 70001, 3 * 320 * 240 * 5 + 7):
        q = rng.integers(0, 256, n, dtype=np.uint8)
        for src, is64 in ((q / 255.0, 1), ((q / 255.0).astype(np.float32), 0)):
            for nt in (1, 3, 0):
                dst = np.full(n, 7, np.uint8)
                assert L.rac_host_pack_gt_u8(src.ctypes.data, is64, n, dst.ctypes.data, nt, ctypes.byref(ok)) == 0
                assert ok.value == 1 and (dst == q).all()
        for bad in (np.nextafter(0.5, 1.0), -1.0 / 255.0, 256.0 / 255.0, np.nan, np.inf, 0.3):
            src = q / 255.0
            src[n // 2] = bad
            assert L.rac_host_pack_gt_u8(src.ctypes.data, 1, n, dst.ctypes.data, 0, ctypes.byref(ok)) == 0
            assert ok.value == 0, (n, bad)
    assert L.rac_host_pack_gt_u8(None, 1, 4, dst.ctypes.data, 0, ctypes.byref(ok)) < 0


def test_host_pack_division_free_check_equals_the_division_form(monkeypatch):
    """The FMA path accepts v iff |255 v - q| < 127.5 ulp(v) (no division); it must agree with the plain
    `(double)q / 255.0 == v` form (RAC_PACK_NO_FMA=1) on every code, on the doubles next to every code, and on tiny /
    signed-zero / out-of-range values."""
    import ctypes
    from rac import _lib
    L = _lib.lib()
    vals = []
    for c in range(256):
        t = c / 255.0
        vals.append(t)
        up = dn = t
        for _ in range(4):
            up, dn = np.nextafter(up, 2.0), np.nextafter(dn, -1.0)
            vals += [up, dn]
    vals += [5e-324, 1e-310, 1e-300, 2.0 ** -900, 2.0 ** -901, 1e-17, -0.0, 1.0 + 2 ** -52, np.nan, np.inf, -np.inf]
    ok = ctypes.c_int(-1)
    dst = np.empty(70000, np.uint8)
    verdicts = {}
    for no_fma in ("0", "1"):
        if no_fma == "1":
            monkeypatch.setenv("RAC_PACK_NO_FMA", "1")
        else:
            monkeypatch.delenv("RAC_PACK_NO_FMA", raising=False)
        out = []
        for v in vals:
            a = np.full(70000, 100 / 255.0)
            a[12345] = v
            assert L.rac_host_pack_gt_u8(a.ctypes.data, 1, a.size, dst.ctypes.data, 2, ctypes.byref(ok)) == 0
            out.append(ok.value)
        verdicts[no_fma] = out
    assert verdicts["0"] == verdicts["1"]
    want = [int(any(v == c / 255.0 for c in range(256))) for v in vals]
    assert verdicts["1"] == want


def test_visualization_overlay_ignores_strangers_and_survives_a_broken_plotting_stack(tmp_path, monkeypatch):
    """The shadow visualization.py overlays the reference's plotting module.  An unrelated visualization.py earlier on
    sys.path must not be executed, and a reference module whose imports fail (matplotlib / cv2 / models missing) must not
    take create_silhouette / blend_imgs down with it."""
    import importlib
    import sys
    import warnings
    stranger = tmp_path / "stranger"
    stranger.mkdir()
    (stranger / "visualization.py").write_text("raise SystemExit('an unrelated visualization.py was executed')\n")
    broken = tmp_path / "broken"
    broken.mkdir()
    (broken / "visualization.py").write_text(
        "import a_plotting_package_that_is_not_installed\n"
        "def create_silhouette(img, silhouette_only=False): pass\n"
        "def blend_imgs(im1, im2, alpha): pass\n"
        "def visualize_examples(model, config, t): pass\n"
        "def create_interpolated_image_fig(a, b, c): pass\n")
    pkg = os.path.join(ROOT, "end-to-end-rgb-pose-estimation-baseline_b200")
    saved = {k: sys.modules.pop(k) for k in ("visualization", "_reference_visualization") if k in sys.modules}
    monkeypatch.setattr(sys, "path", [pkg, str(stranger), str(broken)] +
                        [p for p in sys.path if not os.path.isfile(os.path.join(p or ".", "visualization.py"))])
    try:
        with warnings.catch_warnings(record=True) as w:
            warnings.simplefilter("always")
            viz = importlib.import_module("visualization")
        assert viz.__file__.startswith(pkg)
        assert viz._reference is None and callable(viz.create_silhouette) and callable(viz.blend_imgs)
        assert any("could not be imported" in str(x.message) for x in w)
        assert "_reference_visualization" not in sys.modules
    finally:
        sys.modules.pop("visualization", None)
        sys.modules.update(saved)


def test_mesh_cache_evicts_only_unreferenced_entries_in_batches(monkeypatch):
    """Eviction rules of rac/mesh.py's device cache, on stand-in entries (no GPU): entries somebody outside the cache
    still references survive, the others go least-recently-used first, _EVICT_BATCH at a time."""
    from rac import mesh as M

    class Entry:
        def __init__(self):
            self.destroyed = False
            self.mesh = None

        def destroy(self):
            self.destroyed = True

    monkeypatch.setattr(M, "_cache", type(M._cache)())
    monkeypatch.setattr(M, "MAX_DEVICE_MESHES", 8)
    monkeypatch.setattr(M, "_EVICT_BATCH", 3)
    held = [M._cache_get(k, Entry) for k in ("a", "b")]
    for i in range(20):
        M._cache_get(i, Entry)
        assert len(M._cache) <= 8
    assert [e.destroyed for e in held] == [False, False] and list(M._cache)[:2] == ["a", "b"]
    assert M._cache_get("a", Entry) is held[0] and list(M._cache)[-1] == "a"        # a hit refreshes the entry
    survivors = [k for k in M._cache if isinstance(k, int)]
    assert survivors == sorted(survivors) and survivors[-1] == 19 and len(survivors) <= 6
    # everything pinned: the cache overshoots rather than destroying a mesh in use
    pinned = [M._cache_get(("p", i), Entry) for i in range(12)]
    assert not any(e.destroyed for e in pinned + held)
