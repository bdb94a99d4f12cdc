"""Closed-form known answers for the (?) items of SURVEY.md §8a R5 / R7 / R8 — an independent statement (float64 numpy,
written from the formulas, not from either implementation) that BOTH the CPU oracle (tests/test_oracle_raster.py) and
the CUDA rasteriser (tests/test_gpu_raster.py) are checked against.  `render(mesh, T, K, (W, H)) -> (depth, rgba8)`.

R7 — which BRDF.  pyrender's mesh.frag is the glTF-WebGL-PBR sample shader's metallic-roughness branch:
    D = alpha^2 / (pi ((n.h)^2 (alpha^2 - 1) + 1)^2),  alpha = roughness^2          (Trowbridge-Reitz / GGX; glTF 2.0 App. B)
    G = G1(n.l) G1(n.v),  G1(x) = x / (x (1 - k) + k),  k = (roughness + 1)^2 / 8      (Schlick-GGX, direct-light k)
    F = f0 + (1 - f0)(1 - v.h)^5,  f0 = lerp(0.04, base, metallic)
    f = (1 - F) c_diff / pi + F G D / (4 n.l n.v + 1e-4),  c_diff = base (1 - 0.04)(1 - metallic)
    colour = n.l * radiance * f + base * ambient; n.l, n.v clamped to [0.001, 1]
The glTF 2.0 specification's Appendix B writes the visibility term as the height-correlated Smith-GGX
V = 0.5 / (n.l sqrt(n.v^2 (1 - a^2) + a^2) + n.v sqrt(n.l^2 (1 - a^2) + a^2)) instead of G / (4 n.l n.v); the restatement
follows the SHADER (Schlick k), and `spec_visibility_difference()` bounds what the other choice would change.
"""
import numpy as np


def rx(th):
    c, s = np.cos(th), np.sin(th)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], np.float64)


def quad(size, material=None, uv=False, tex=None):
    from rac import mesh as M
    v = np.array([[-size, -size, 0], [size, -size, 0], [size, size, 0], [-size, size, 0]], np.float32)
    f = np.array([[0, 2, 1], [0, 3, 2]], np.int32)          # normal -z: faces a camera at the origin when placed at +z
    t = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32) if uv else None
    return M.Mesh(v, f, uv=t, tex=tex, material=None if material is None else np.asarray(material, np.float32))


def brdf_colour(p, n, base, metallic, rough, ambient=0.1, intensity=15.0, visibility="schlick"):
    """8-bit colour of a fragment at camera-frame point p with unit normal n under the head-light spot (renderer.py:50-57)."""
    dist = np.linalg.norm(p)
    l = -p / dist                       # light = camera = origin: l = v = h
    ndl = float(n @ l)
    nl = min(max(ndl, 0.001), 1.0); nv = min(max(abs(ndl), 0.001), 1.0); nh = min(max(ndl, 0.0), 1.0)
    f0 = 0.04 + (base - 0.04) * metallic            # F = f0: v.h = 1
    cdiff = base * 0.96 * (1.0 - metallic)
    a = min(max(rough, 0.04), 1.0) ** 2
    a2 = a * a
    D = a2 / (np.pi * (nh * nh * (a2 - 1.0) + 1.0) ** 2)
    if visibility == "schlick":
        k = (rough + 1.0) ** 2 / 8.0
        G = (nl / (nl * (1 - k) + k)) * (nv / (nv * (1 - k) + k))
        spec = f0 * G * D / (4 * nl * nv + 1e-4)
    else:                               # glTF 2.0 Appendix B: height-correlated Smith-GGX visibility
        V = 0.5 / (nl * np.sqrt(nv * nv * (1 - a2) + a2) + nv * np.sqrt(nl * nl * (1 - a2) + a2))
        spec = f0 * V * D
    diff = (1.0 - f0) * cdiff / np.pi
    las = 1.0 / max(0.001, np.cos(np.pi / 9) - np.cos(np.pi / 2))           # pyrender SpotLight uniforms
    lao = -np.cos(np.pi / 2) * las
    cd = min(max(p[2] / dist, 0.0), 1.0)
    att = min(max(cd * las + lao, 0.0), 1.0) ** 2 / dist ** 2
    col = nl * att * intensity * (diff + spec) + base * ambient
    return int(min(max(col, 0.0) ** (1 / 2.2), 1.0) * 255 + 0.5)


BRDF_POINTS = [   # (tilt of the quad about x in degrees -> n.l = cos(tilt) at the image centre, roughness, metallic, base)
    (0.0, 0.8, 0.2, 0.3),        # pyrender's default material, head on
    (40.0, 0.8, 0.2, 0.3),       # ... at n.l = 0.77
    (25.0, 0.35, 0.9, 0.7),      # smooth metal: the specular term dominates
    (60.0, 0.5, 0.0, 0.9),       # dielectric at a grazing-ish angle
]


def check_brdf_points(render, tol=1):
    """Tilted quads: the colour at and around the image centre equals the closed form."""
    W, H, d = 96, 96, 2.0
    fx = 50.0 / (36.0 / W)
    K = np.array([[[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]]], np.float32)
    for tilt, rough, metallic, base in BRDF_POINTS:
        R = rx(np.deg2rad(tilt))
        T = np.eye(4, dtype=np.float32)[None].copy()
        T[0, :3, :3] = R.astype(np.float32)
        T[0, 2, 3] = d
        n = R @ np.array([0.0, 0.0, -1.0])
        depth, rgba8 = render(quad(0.3, (base, base, base, 1.0, metallic, rough)), T, K, (W, H))
        for (r, c) in [(H // 2, W // 2), (H // 2 - 3, W // 2 + 4), (H // 2 + 5, W // 2 - 6)]:
            ray = np.array([(c + 0.5 - W / 2.0) / fx, (r + 0.5 - H / 2.0) / fx, 1.0])
            t = (n @ np.array([0.0, 0.0, d])) / (n @ ray)
            p = t * ray
            want = brdf_colour(p, n, base, metallic, rough)
            got = rgba8[0, r, c, :3].astype(int)
            ray0 = np.array([(c + 96.0 / 256.0 - W / 2.0) / fx, (r + 32.0 / 256.0 - H / 2.0) / fx, 1.0])      # depth = sample 0 (R5)
            z0 = (n @ np.array([0.0, 0.0, d])) / (n @ ray0)
            assert depth[0, r, c] > 0 and abs(float(depth[0, r, c]) - z0) < 2e-5 * z0, (tilt, r, c, depth[0, r, c], z0)
            assert (np.abs(got - want) <= tol).all(), (tilt, rough, metallic, base, r, c, got, want)


def spec_visibility_difference():
    """Largest 8-bit colour difference between the shader's Schlick-k visibility and the glTF spec's Smith-GGX one over a
    grid of head-light configurations at 2.5 m (DESIGN.md records the number)."""
    worst = 0
    for tilt in np.linspace(0, 80, 17):
        n = rx(np.deg2rad(tilt)) @ np.array([0.0, 0.0, -1.0])
        for rough, metallic, base in [(0.8, 0.2, 0.3), (0.35, 0.9, 0.7), (0.5, 0.0, 0.9), (1.0, 1.0, 1.0), (0.2, 0.5, 0.5)]:
            p = np.array([0.0, 0.0, 2.5])
            worst = max(worst, abs(brdf_colour(p, n, base, metallic, rough) - brdf_colour(p, n, base, metallic, rough, visibility="smith")))
    return worst


def check_half_covered_pixels(render):
    """R5: colour = (sum of the 4 samples' colours + 2) >> 2 with background 0; depth = SAMPLE 0 only.  A fronto-parallel
    quad whose right edge runs through a pixel column at x = c + 0.5 covers the samples at x-offsets 0.375 and 0.125
    (samples 0 and 2) -> half the colour, depth present; with the edge at c + 0.25 only sample 2 (0.125) is inside ->
    a quarter of the colour and NO depth."""
    W, H, d, fx = 64, 48, 2.0, 100.0
    K = np.array([[[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]]], np.float32)
    from rac import mesh as M
    for frac, n_cov, has_depth in [(0.5, 2, True), (0.25, 1, False), (0.75, 3, True), (0.9, 4, True)]:
        c = 40
        x_right = (c + frac - W / 2.0) * d / fx
        v = np.array([[-0.2, -0.15, 0], [x_right, -0.15, 0], [x_right, 0.15, 0], [-0.2, 0.15, 0]], np.float32)
        m = M.Mesh(v, np.array([[0, 2, 1], [0, 3, 2]], np.int32))
        T = np.eye(4, dtype=np.float32)[None].copy()
        T[0, 2, 3] = d
        depth, rgba8 = render(m, T, K, (W, H))
        r = H // 2
        full = rgba8[0, r, c - 3, :3].astype(int)                 # interior pixel, same row (colour varies slowly: compare loosely)
        edge = rgba8[0, r, c].astype(int)
        assert rgba8[0, r, c - 3, 3] == 255 and depth[0, r, c - 3] > 0
        assert edge[3] == (255 * n_cov + 2) >> 2, (frac, edge)
        assert (np.abs(edge[:3] - ((full * n_cov + 2) >> 2)) <= 2).all(), (frac, edge, full)
        assert (depth[0, r, c] > 0) == has_depth, (frac, depth[0, r, c])
        assert rgba8[0, r, c + 1, 3] == 0 and depth[0, r, c + 1] == 0


def check_minified_checker(render, tol=1):
    """R8: trilinear filtering on a mip chain built from the ENCODED texels.  A 64x64 texture alternating 0 / 255 every
    texel, mapped onto a fronto-parallel quad that covers 16x16 pixels (4 texels per pixel -> LOD 2): level 1 and below
    are the flat average (0 + 255 + 0 + 255 + 2) >> 2 = 128, so every interior pixel must show base = srgb_to_linear(128/255)
    — nearest / bilinear sampling of level 0 would give a moire of blacks and whites instead."""
    W = H = 64
    fx, d = 100.0, 2.0
    K = np.array([[[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]]], np.float32)
    yy, xx = np.mgrid[0:64, 0:64]
    tex = np.repeat((((xx + yy) % 2) * 255).astype(np.uint8)[..., None], 3, axis=2)
    size = 8.0 * d / fx                  # half-size: 16 px across
    m = quad(size, (1.0, 1.0, 1.0, 1.0, 0.2, 0.8), uv=True, tex=tex)
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, 2, 3] = d
    depth, rgba8 = render(m, T, K, (W, H))
    s = 128.0 / 255.0
    base = ((s + 0.055) / 1.055) ** 2.4
    for (r, c) in [(H // 2, W // 2), (H // 2 - 4, W // 2 + 3), (H // 2 + 5, W // 2 - 5)]:
        p = np.array([(c + 0.5 - W / 2.0) / fx * d, (r + 0.5 - H / 2.0) / fx * d, d])
        want = brdf_colour(p, np.array([0.0, 0.0, -1.0]), base, 0.2, 0.8)
        got = rgba8[0, r, c, :3].astype(int)
        assert (np.abs(got - want) <= tol).all(), (r, c, got, want)
    inner = rgba8[0, H // 2 - 6:H // 2 + 6, W // 2 - 6:W // 2 + 6, 0].astype(int)
    assert inner.max() - inner.min() <= 3, "a minified 1-texel checker must resolve to a flat grey"


def check_near_plane_clips_per_sample(render):
    """GL clips a triangle at z = znear (0.05 m); the spec rejects every sample with z < znear in the depth test instead,
    which selects the same samples.  A long floor-like quad running from z = 0.02 m to z = 1.0 m below the optical axis:
    every covered pixel has depth >= znear, coverage reaches right down to the plane, and the rows that would show the
    part in front of the plane stay empty."""
    from rac import mesh as M
    W, H, fx = 96, 96, 100.0
    K = np.array([[[fx, 0, W / 2.0], [0, fx, H / 2.0], [0, 0, 1]]], np.float32)
    y0 = 0.02                                   # 2 cm below the axis (image y grows downwards)
    v = np.array([[-0.01, y0, 0.02], [0.01, y0, 0.02], [0.3, y0, 1.0], [-0.3, y0, 1.0]], np.float32)
    f = np.array([[0, 1, 2], [0, 2, 3]], np.int32)      # seen from above (camera at the origin, looking along +z)
    T = np.eye(4, dtype=np.float32)[None].copy()
    depth, rgba8 = render(M.Mesh(v, f), T, K, (W, H))
    cov = depth[0] > 0
    assert cov.sum() > 50, "the part of the floor beyond the near plane must be visible"
    assert depth[0][cov].min() >= 0.05 - 1e-6
    assert depth[0][cov].min() < 0.0515, "coverage must reach down to the near plane (depth %.4f)" % depth[0][cov].min()
    # a point at depth z projects to row cy + fx * y0 / z: z = znear -> row 88; nothing may be drawn below it
    row_near = H / 2.0 + fx * y0 / 0.05
    rows = np.nonzero(cov.any(1))[0]
    assert rows.max() <= int(row_near) and rows.max() >= int(row_near) - 1
    assert rows.min() <= H // 2 + 3              # the far end (z = 1 m) sits 2 px below the centre row
