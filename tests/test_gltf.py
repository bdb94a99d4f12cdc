"""Binary glTF ingestion (rac/gltf.py) — the loader half of renderer.py:28-48 for the reference's "texturized" meshes:
every triangle primitive of the file, in its local frame (node transforms ignored, as `scene.geometry.values()` does),
rotated by Rx(90 deg), with the file's pbrMetallicRoughness material.  CPU only (no device calls)."""
import json
import struct

import numpy as np
import pytest

from rac import gltf
from rac import mesh as M


def _prims():
    a = M.with_texture(M.icosphere(2), size=32, seed=2)
    b = M.box(0.5, 0.3, 0.2, 1)
    vc = np.tile(np.array([[10, 200, 30, 255]], np.uint8), (len(b.verts), 1))
    return [dict(verts=a.verts, faces=a.faces, uv=a.uv, tex=a.tex, material=(1, 1, 1, 1, 0.1, 0.7)),
            dict(verts=b.verts + np.float32([0.6, 0, 0]), faces=b.faces, material=(0.8, 0.2, 0.1, 1, 0.5, 0.4), vcolor=vc),
            dict(verts=b.verts + np.float32([-0.6, 0, 0]), faces=b.faces, vcolor=vc)]


def test_roundtrip_matches_the_reference_semantics(tmp_path):
    prims = _prims()
    path = str(tmp_path / "scene.glb")
    gltf.save_glb(path, prims, node_translation=(5, 5, 5))          # a faithful loader ignores the node transform
    m = M.load_mesh(path)
    assert len(m.parts) == 3 and m.face_part.dtype == np.uint8
    nv = np.cumsum([0] + [len(p["verts"]) for p in prims])
    nf = np.cumsum([0] + [len(p["faces"]) for p in prims])
    for k, p in enumerate(prims):
        v = np.asarray(p["verts"], np.float64)
        want = np.stack([v[:, 0], -v[:, 2], v[:, 1]], 1)             # Rx(90 deg), renderer.py:38-40
        np.testing.assert_allclose(m.verts[nv[k]:nv[k + 1]], want, atol=1e-7)
        np.testing.assert_array_equal(m.faces[nf[k]:nf[k + 1]], np.asarray(p["faces"]) + nv[k])
        assert (m.face_part[nf[k]:nf[k + 1]] == k).all()
    # materials: the file's own pbrMetallicRoughness, baseColorFactor in 1/255 steps (trimesh keeps it as uint8)
    np.testing.assert_allclose(m.parts[0]["material"], (1, 1, 1, 1, 0.1, 0.7), atol=1e-7)
    np.testing.assert_allclose(m.parts[1]["material"], (204 / 255, 51 / 255, 26 / 255, 1, 0.5, 0.4), atol=1e-7)
    np.testing.assert_array_equal(m.parts[0]["tex"], prims[0]["tex"])
    np.testing.assert_allclose(m.uv[nv[0]:nv[1]], prims[0]["uv"], atol=1e-7)
    # COLOR_0: ignored on a primitive WITH a material, used (on pyrender's white vertex-colour material) on one WITHOUT
    assert not m.parts[1]["use_vcolor"] and (m.vcolor[nv[1]:nv[2]] == 255).all()
    assert m.parts[2]["use_vcolor"] and (m.vcolor[nv[2]:nv[3]] == [10, 200, 30, 255]).all()
    np.testing.assert_allclose(m.parts[2]["material"], (1.0, 1.0, 1.0, 1.0, 0.2, 0.8))
    assert m.parts[1]["tex"] is None and m.parts[2]["tex"] is None
    # the trimesh-compatible vertex source of data_loaders.get_vertices
    assert m.vertices is m.verts


def test_gltf_defaults_strides_and_errors(tmp_path):
    """Hand-built file: interleaved (strided) POSITION + TEXCOORD_0 view, uint16 indices, a material that omits every
    factor (glTF defaults: white, metallic 1, roughness 1), a non-triangle primitive that must be skipped."""
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], "<f4")
    uv = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], "<f4")
    inter = np.concatenate([pos, uv], 1).astype("<f4").tobytes()      # stride 20
    idx = np.array([0, 1, 2, 0, 2, 3], "<u2").tobytes()
    blob = inter + idx
    g = {"asset": {"version": "2.0"}, "buffers": [{"byteLength": len(blob)}],
         "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": len(inter), "byteStride": 20},
                         {"buffer": 0, "byteOffset": len(inter), "byteLength": len(idx)}],
         "accessors": [{"bufferView": 0, "byteOffset": 0, "componentType": 5126, "count": 4, "type": "VEC3"},
                       {"bufferView": 0, "byteOffset": 12, "componentType": 5126, "count": 4, "type": "VEC2"},
                       {"bufferView": 1, "componentType": 5123, "count": 6, "type": "SCALAR"}],
         "materials": [{}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "TEXCOORD_0": 1}, "indices": 2, "material": 0},
                                    {"attributes": {"POSITION": 0}, "indices": 2, "mode": 1}]}]}
    js = json.dumps(g).encode()
    js += b" " * (-len(js) % 4)
    blob += b"\0" * (-len(blob) % 4)
    path = str(tmp_path / "hand.glb")
    with open(path, "wb") as f:
        f.write(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(blob)))
        f.write(struct.pack("<I4s", len(js), b"JSON") + js + struct.pack("<I4s", len(blob), b"BIN\0") + blob)
    m = gltf.load_glb(path)
    assert len(m.parts) == 1 and len(m.faces) == 2
    np.testing.assert_allclose(m.parts[0]["material"], (1, 1, 1, 1, 1, 1))
    np.testing.assert_allclose(m.verts, [[0, 0, 0], [1, 0, 0], [0, 0, 1], [0, -1, 0]], atol=1e-7)
    np.testing.assert_allclose(m.uv, uv)
    raw = gltf.load_glb(path, apply_rx90=False)
    np.testing.assert_allclose(raw.verts, pos)
    bad = str(tmp_path / "bad.glb")
    with open(bad, "wb") as f:
        f.write(b"glTX" + b"\0" * 20)
    with pytest.raises(ValueError):
        gltf.load_glb(bad)
    with pytest.raises(ValueError):
        M.load_mesh(str(tmp_path / "x.obj"))


def test_oracle_renders_every_part_with_its_own_material(tmp_path):
    """The CPU oracle with per-face parts: each primitive shows its own colour (textured / red factor / green COLOR_0)."""
    from conftest import intrinsics
    from oracle import raster_oracle as ro
    path = str(tmp_path / "scene.glb")
    gltf.save_glb(path, _prims())
    m = M.load_mesh(path)
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, :3, 3] = [0.0, 0.0, 2.5]
    K = intrinsics(1, 200, 200)
    _, depth, q = ro.render_batch(m, T, K, 200)
    assert (depth > 0).mean() > 0.05
    cov = q[0][depth[0] > 0][:, :3].astype(int)
    red = (cov[:, 0] > cov[:, 1] + 50) & (cov[:, 0] > cov[:, 2] + 50)
    green = (cov[:, 1] > cov[:, 0] + 50) & (cov[:, 1] > cov[:, 2] + 50)
    assert red.sum() > 100 and green.sum() > 100 and (~red & ~green).sum() > 100


def test_vertex_coloured_mesh_without_material_gets_white_base(tmp_path):
    """pyrender's material selection (Mesh.from_trimesh -> _get_trimesh_props, reached from renderer.py:41): a trimesh
    whose visual kind is 'vertex' and that has no material is drawn with baseColorFactor (1,1,1,1), metallic 0.2,
    roughness 0.8 — the 0.3 grey is for meshes with no colours at all.  Known answer: a mesh with a UNIFORM vertex
    colour c must render exactly like the uncoloured mesh with base colour c / 255, for coloured PLY files and for
    material-less COLOR_0 primitives of a .glb alike."""
    from conftest import intrinsics
    from oracle import raster_oracle as ro
    s = M.icosphere(2)
    c = np.array([200, 120, 40, 255], np.uint8)
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, :3, 3] = [0.05, -0.02, 2.5]
    K = intrinsics(1, 96, 96)
    plain = M.Mesh(s.verts, s.faces, material=(200 / 255, 120 / 255, 40 / 255, 1.0, 0.2, 0.8))
    _, d_ref, q_ref = ro.render_batch(plain, T, K, 96)
    assert (d_ref > 0).mean() > 0.1
    # PLY with per-vertex colours, loaded back from disk
    ply = str(tmp_path / "c.ply")
    M.save_ply(ply, M.Mesh(s.verts, s.faces, vcolor=np.tile(c, (len(s.verts), 1))))
    m = M.load_mesh(ply)
    assert m.material is None and list(m.material_array()) == pytest.approx([1, 1, 1, 1, 0.2, 0.8])
    assert list(M.Mesh(s.verts, s.faces).material_array()) == pytest.approx([0.3, 0.3, 0.3, 1, 0.2, 0.8])
    _, d1, q1 = ro.render_batch(m, T, K, 96)
    np.testing.assert_array_equal(d1, d_ref)
    assert np.abs(q1.astype(int) - q_ref.astype(int)).max() <= 1
    # .glb: COLOR_0 on a primitive without a material (loader applies Rx(90 deg): pre-rotate the vertices back)
    v = np.stack([s.verts[:, 0], s.verts[:, 2], -s.verts[:, 1]], 1)
    glb = str(tmp_path / "c.glb")
    gltf.save_glb(glb, [dict(verts=v, faces=s.faces, vcolor=np.tile(c, (len(v), 1)))])
    g = M.load_mesh(glb)
    assert g.parts[0]["use_vcolor"] and list(g.parts[0]["material"]) == pytest.approx([1, 1, 1, 1, 0.2, 0.8])
    _, d2, q2 = ro.render_batch(g, T, K, 96)
    assert ((d2 > 0) != (d_ref > 0)).mean() <= 1e-3
    both = (d2[0] > 0) & (d_ref[0] > 0)
    assert np.abs(q2[0][both].astype(int) - q_ref[0][both].astype(int)).max() <= 1


def test_double_sided_material_draws_back_faces_unlit(tmp_path):
    """material.doubleSided (Blender's default export): pyrender disables back-face culling for that material and its
    shader does not flip the normal, so a quad seen from behind is covered (alpha 255, depth present) and shows ambient
    light only; with doubleSided false it is culled."""
    from conftest import intrinsics
    from oracle import raster_oracle as ro
    import known_answers as KA
    q = KA.quad(0.3)
    T = np.eye(4, dtype=np.float32)[None].copy()
    T[0, :3, :3] = np.diag([1.0, -1.0, -1.0])           # flip the quad: the camera looks at its back
    T[0, 2, 3] = 2.0
    K = intrinsics(1, 96, 96)
    out = {}
    for ds in (False, True):
        path = str(tmp_path / ("quad_%d.glb" % ds))
        # (the loader's Rx(90) is undone by pre-rotating the vertices: (x, y, z) -> (x, z, -y))
        v = np.stack([q.verts[:, 0], q.verts[:, 2], -q.verts[:, 1]], 1)
        gltf.save_glb(path, [dict(verts=v, faces=q.faces, material=(0.8, 0.8, 0.8, 1.0, 0.0, 0.5), double_sided=ds)])
        m = M.load_mesh(path)
        assert m.parts[0]["double_sided"] is ds
        np.testing.assert_allclose(m.verts, q.verts, atol=1e-7)
        _, depth, rgba = ro.render_batch(m, T, K, 96)
        out[ds] = (depth, rgba)
    assert (out[False][0] == 0).all() and (out[False][1][..., 3] == 0).all()
    depth, rgba = out[True]
    c = rgba[0, 48, 48]
    assert depth[0, 48, 48] == pytest.approx(2.0, rel=1e-4) and c[3] == 255
    # ambient only: base * 0.1 (+ n.l clamped to 0.001 of direct light) -> (0.8 * 0.1)^(1/2.2) * 255 = 81
    assert abs(int(c[0]) - 81) <= 2 and c[0] == c[1] == c[2]
    # the front of the same quad is lit
    T[0, :3, :3] = np.eye(3)
    _, _, front = ro.render_batch(m, T, K, 96)
    assert front[0, 48, 48, 0] > 150


def _write_glb(path, g, blob):
    js = json.dumps(g).encode()
    js += b" " * (-len(js) % 4)
    blob += b"\0" * (-len(blob) % 4)
    with open(path, "wb") as f:
        f.write(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(blob)))
        f.write(struct.pack("<I4s", len(js), b"JSON") + js + struct.pack("<I4s", len(blob), b"BIN\0") + blob)


def test_accessor_variants(tmp_path):
    """Un-indexed primitive (POSITION order = triangle list), float VEC3 COLOR_0 on a material-less primitive, normalised
    uint16 TEXCOORD_0, uint8 indices, an accessor without a bufferView (zeros), and a file with no triangle primitive."""
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], "<f4")
    col = np.array([[1.0, 0.5, 0.0]] * 6, "<f4")
    uv16 = np.array([[0, 0], [65535, 0], [0, 65535], [65535, 0], [65535, 65535], [0, 65535]], "<u2")
    idx8 = np.array([0, 1, 2, 3, 4, 5], "u1")
    parts = [pos.tobytes(), col.tobytes(), uv16.tobytes(), idx8.tobytes()]
    offs = np.cumsum([0] + [len(p) + (-len(p) % 4) for p in parts])
    blob = b"".join(p + b"\0" * (-len(p) % 4) for p in parts)
    views = [{"buffer": 0, "byteOffset": int(offs[i]), "byteLength": len(parts[i])} for i in range(4)]
    g = {"asset": {"version": "2.0"}, "buffers": [{"byteLength": len(blob)}], "bufferViews": views,
         "accessors": [{"bufferView": 0, "componentType": 5126, "count": 6, "type": "VEC3"},
                       {"bufferView": 1, "componentType": 5126, "count": 6, "type": "VEC3"},
                       {"bufferView": 2, "componentType": 5123, "count": 6, "type": "VEC2", "normalized": True},
                       {"bufferView": 3, "componentType": 5121, "count": 6, "type": "SCALAR"},
                       {"componentType": 5126, "count": 6, "type": "VEC2"}],
         "materials": [{"pbrMetallicRoughness": {"baseColorFactor": [0.5, 0.25, 1.0, 1.0]}, "doubleSided": True}],
         "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "COLOR_0": 1}},                          # no indices, no material
                                    {"attributes": {"POSITION": 0, "TEXCOORD_0": 2}, "indices": 3, "material": 0},
                                    {"attributes": {"POSITION": 0, "TEXCOORD_0": 4}, "material": 0}]}]}
    path = str(tmp_path / "variants.glb")
    _write_glb(path, g, blob)
    m = gltf.load_glb(path, apply_rx90=False)
    assert len(m.parts) == 3 and len(m.faces) == 6 and len(m.verts) == 18
    np.testing.assert_array_equal(m.faces[:2], [[0, 1, 2], [3, 4, 5]])
    assert m.parts[0]["use_vcolor"] and (m.vcolor[:6] == [255, 128, 0, 255]).all()          # float VEC3 colours, alpha 255
    np.testing.assert_allclose(m.uv[6:12], uv16 / 65535.0, atol=1e-7)                       # normalised u16 texcoords
    np.testing.assert_array_equal(m.faces[2:4] - 6, [[0, 1, 2], [3, 4, 5]])                 # u8 indices
    assert (m.uv[12:18] == 0).all()                                                         # accessor without a bufferView
    np.testing.assert_allclose(m.parts[1]["material"], (128 / 255, 64 / 255, 1.0, 1.0, 1.0, 1.0))
    assert m.parts[1]["double_sided"] and not m.parts[0]["double_sided"]
    g2 = dict(g, meshes=[{"primitives": [{"attributes": {"POSITION": 0}, "mode": 0}]}])
    p2 = str(tmp_path / "points.glb")
    _write_glb(p2, g2, blob)
    with pytest.raises(ValueError):
        gltf.load_glb(p2)
