"""Binary glTF (.glb) ingestion — the reference's "texturized" meshes (old-configs/effnet_b3_svd.py:15,
old-configs/effnetv2_m_normdepth.py:38: model3d_dataset = "ModelNet10-texturized").

What the reference does with such a file (renderer.py:28-48):
    trimesh_mesh = tm.load(path)                       -> a trimesh.Scene for .glb
    for m in trimesh_mesh.geometry.values():           every geometry (= glTF mesh primitive) of the file, in its OWN
        m.apply_transform(Rx(90 deg))                  local frame — the scene graph's node transforms are NOT applied —
        scene.add(pyrender.Mesh.from_trimesh(m, smooth=False))      rotated by Rx(90) ("correction for glb export from
                                                       blender"), flat-shaded, with the material the file gives it
so all primitives land in one pyrender scene and one depth buffer.  load_glb() returns the same thing as ONE merged,
indexed mesh with a part index per face (rac_mesh_create_parts): geometry identical, one material / texture per part.

Material mapping (trimesh PBRMaterial -> pyrender MetallicRoughnessMaterial, pyrender/mesh.py _get_trimesh_props):
    baseColorFactor (default 1,1,1,1; trimesh stores it as uint8 RGBA, so it is rounded to 1/255 steps here too),
    metallicFactor, roughnessFactor (glTF / pyrender defaults 1.0, 1.0), baseColorTexture (TEXCOORD_0, GL_REPEAT, mips);
    a primitive WITHOUT a material gets pyrender's default material (0.3 grey, metallic 0.2, roughness 0.8) and, if it
    carries COLOR_0, vertex colours; COLOR_0 of a primitive WITH a material is ignored (trimesh parks it in
    vertex_attributes, pyrender never sees it).
    material.doubleSided (what Blender writes unless "backface culling" is ticked): the part's back faces are drawn too —
    pyrender switches GL_CULL_FACE off for such a material and its mesh.frag does not flip the normal, so they receive
    ambient light only.
Not applied (documented in DESIGN.md §7): normal / occlusion / emissive / metallicRoughness textures, alpha blending,
KHR extensions, sparse accessors, non-triangle primitives.
No third-party parser: struct + json + PIL for the embedded PNG / JPEG images.
"""
import io
import json
import struct

import numpy as np

_COMP = {5120: ("i1", 1), 5121: ("u1", 1), 5122: ("i2", 2), 5123: ("u2", 2), 5125: ("u4", 4), 5126: ("f4", 4)}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT2": 4, "MAT3": 9, "MAT4": 16}
DEFAULT_MATERIAL = (0.3, 0.3, 0.3, 1.0, 0.2, 0.8)          # pyrender's default for primitives without a material
VERTEX_COLOR_MATERIAL = (1.0, 1.0, 1.0, 1.0, 0.2, 0.8)     # ... and without a material but WITH COLOR_0: white base, the vertex
                                                            # colour is the base colour (pyrender _get_trimesh_props, kind 'vertex')


def _chunks(data):
    magic, version, length = struct.unpack_from("<4sII", data, 0)
    if magic != b"glTF" or version != 2:
        raise ValueError("not a binary glTF 2.0 file")
    off, out = 12, {}
    while off + 8 <= min(length, len(data)):
        clen, ctype = struct.unpack_from("<I4s", data, off)
        out.setdefault(ctype, data[off + 8: off + 8 + clen])
        off += 8 + clen + (-clen % 4)
    if b"JSON" not in out:
        raise ValueError("glb without a JSON chunk")
    return json.loads(out[b"JSON"].decode("utf-8")), out.get(b"BIN\x00", b"")


def _accessor(gltf, binary, idx):
    acc = gltf["accessors"][idx]
    if "sparse" in acc:
        raise ValueError("sparse accessors are not supported")
    dt, size = _COMP[acc["componentType"]]
    nc = _NCOMP[acc["type"]]
    count = acc["count"]
    if "bufferView" not in acc:
        return np.zeros((count, nc), dtype=dt)
    bv = gltf["bufferViews"][acc["bufferView"]]
    if bv.get("buffer", 0) != 0:
        raise ValueError("only the embedded BIN buffer is supported")
    base = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
    stride = bv.get("byteStride", 0) or nc * size
    if stride == nc * size:
        arr = np.frombuffer(binary, dtype="<" + dt, count=count * nc, offset=base).reshape(count, nc)
    else:
        raw = np.frombuffer(binary, dtype=np.uint8, count=(count - 1) * stride + nc * size, offset=base)
        idxs = (np.arange(count)[:, None] * stride + np.arange(nc * size)[None, :]).reshape(-1)
        arr = raw[idxs].view("<" + dt).reshape(count, nc)
    if acc.get("normalized") and dt != "f4":
        scale = {"u1": 255.0, "u2": 65535.0, "i1": 127.0, "i2": 32767.0}[dt]
        arr = np.maximum(arr.astype(np.float32) / scale, -1.0)
    return arr


def _image_rgb(gltf, binary, tex_index):
    from PIL import Image
    tex = gltf["textures"][tex_index]
    img = gltf["images"][tex["source"]]
    if "bufferView" not in img:
        raise ValueError("only images embedded in the BIN chunk are supported")
    bv = gltf["bufferViews"][img["bufferView"]]
    raw = binary[bv.get("byteOffset", 0): bv.get("byteOffset", 0) + bv["byteLength"]]
    return np.asarray(Image.open(io.BytesIO(raw)).convert("RGB"), dtype=np.uint8)      # row 0 = top = glTF v 0


def _material(gltf, binary, idx):
    """-> dict(material=(r,g,b,a,metallic,roughness), tex=RGB8 or None)."""
    m = gltf["materials"][idx]
    pbr = m.get("pbrMetallicRoughness", {})
    bc = np.asarray(pbr.get("baseColorFactor", [1.0, 1.0, 1.0, 1.0]), dtype=np.float64)
    bc = np.rint(np.clip(bc, 0.0, 1.0) * 255.0) / 255.0                    # trimesh keeps the factor as uint8 RGBA
    tex = None
    if "baseColorTexture" in pbr:
        tex = _image_rgb(gltf, binary, pbr["baseColorTexture"]["index"])
    return {"material": (bc[0], bc[1], bc[2], bc[3], float(pbr.get("metallicFactor", 1.0)), float(pbr.get("roughnessFactor", 1.0))),
            "tex": tex, "double_sided": bool(m.get("doubleSided", False))}


def load_glb(path, apply_rx90=True):
    """-> rac.mesh.Mesh with .face_part (T,) u8 and .parts [dict(material, tex, use_vcolor)] (one entry per triangle
    primitive of the file, in file order).  apply_rx90: the reference's correction (renderer.py:38-40)."""
    from .mesh import Mesh
    with open(path, "rb") as f:
        gltf, binary = _chunks(f.read())
    V, F, UV, VC, FP, parts = [], [], [], [], [], []
    nv = 0
    mat_cache = {}
    for mesh in gltf.get("meshes", []):
        for prim in mesh.get("primitives", []):
            if prim.get("mode", 4) != 4 or "POSITION" not in prim.get("attributes", {}):
                continue            # points / lines: trimesh loads them as non-mesh geometry, pyrender would not draw triangles
            attr = prim["attributes"]
            pos = _accessor(gltf, binary, attr["POSITION"]).astype(np.float64)
            if "indices" in prim:
                idx = _accessor(gltf, binary, prim["indices"]).astype(np.int64).reshape(-1)
            else:
                idx = np.arange(len(pos), dtype=np.int64)
            idx = idx[: len(idx) // 3 * 3].reshape(-1, 3)
            if apply_rx90:          # Rx(90 deg): (x, y, z) -> (x, -z, y)
                pos = np.stack([pos[:, 0], -pos[:, 2], pos[:, 1]], axis=1)
            has_mat = "material" in prim and "materials" in gltf
            if has_mat:
                if prim["material"] not in mat_cache:
                    mat_cache[prim["material"]] = _material(gltf, binary, prim["material"])
                part = dict(mat_cache[prim["material"]], use_vcolor=False)
            else:
                has_vc = "COLOR_0" in attr
                part = {"material": VERTEX_COLOR_MATERIAL if has_vc else DEFAULT_MATERIAL, "tex": None, "use_vcolor": has_vc,
                        "double_sided": False}
            uv = np.zeros((len(pos), 2), dtype=np.float32)
            if has_mat and "TEXCOORD_0" in attr:
                uv = _accessor(gltf, binary, attr["TEXCOORD_0"]).astype(np.float32)[:, :2]
            elif part["tex"] is not None:
                part = dict(part, tex=None)                   # a texture without texture coordinates cannot be sampled
            vc = np.full((len(pos), 4), 255, dtype=np.uint8)
            if part["use_vcolor"]:
                c = np.asarray(_accessor(gltf, binary, attr["COLOR_0"]))
                if c.dtype.kind == "f":                       # float or normalised integer colours in [0, 1]
                    c = np.rint(np.clip(c, 0.0, 1.0) * 255.0)
                vc[:, : c.shape[1]] = c.astype(np.uint8)
            V.append(pos); F.append(idx + nv); UV.append(uv); VC.append(vc)
            FP.append(np.full(len(idx), len(parts), dtype=np.uint8))
            parts.append(part)
            nv += len(pos)
    if not parts:
        raise ValueError("no triangle primitives in " + str(path))
    if len(parts) > 255:
        raise ValueError("more than 255 primitives in " + str(path))
    mesh = Mesh(np.concatenate(V), np.concatenate(F), uv=np.concatenate(UV), vcolor=np.concatenate(VC))
    mesh.face_part = np.ascontiguousarray(np.concatenate(FP))
    mesh.parts = parts
    return mesh


# ---- writer (tests and synthetic "texturized" datasets) ------------------------------------------------------------
def save_glb(path, prims, node_translation=None):
    """prims: list of dict(verts (V,3), faces (T,3), uv (V,2)?, vcolor (V,4) u8?, material (6,)? , tex RGB8 (h,w,3)?,
    double_sided bool?).
    One glTF mesh with one primitive per entry, each with its own material when `material` / `tex` is given; images are
    embedded as PNG.  node_translation: a node transform that a faithful loader must IGNORE (the reference iterates
    scene.geometry, not the scene graph)."""
    from PIL import Image
    blob = bytearray()
    views, accessors, materials, textures, images, gprims = [], [], [], [], [], []

    def add_view(data, target=None):
        while len(blob) % 4:
            blob.append(0)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)}
        if target:
            v["target"] = target
        blob.extend(data)
        views.append(v)
        return len(views) - 1

    def add_acc(arr, ctype, typ, target=None, normalized=False, minmax=False):
        a = {"bufferView": add_view(arr.tobytes(), target), "componentType": ctype, "count": len(arr), "type": typ}
        if normalized:
            a["normalized"] = True
        if minmax:
            a["min"], a["max"] = arr.min(0).tolist(), arr.max(0).tolist()
        accessors.append(a)
        return len(accessors) - 1

    for p in prims:
        verts = np.ascontiguousarray(p["verts"], dtype="<f4")
        faces = np.ascontiguousarray(p["faces"], dtype="<u4").reshape(-1)
        attrs = {"POSITION": add_acc(verts, 5126, "VEC3", 34962, minmax=True)}
        if p.get("uv") is not None:
            attrs["TEXCOORD_0"] = add_acc(np.ascontiguousarray(p["uv"], dtype="<f4"), 5126, "VEC2", 34962)
        if p.get("vcolor") is not None:
            attrs["COLOR_0"] = add_acc(np.ascontiguousarray(p["vcolor"], dtype=np.uint8), 5121, "VEC4", 34962, normalized=True)
        prim = {"attributes": attrs, "indices": add_acc(faces, 5125, "SCALAR", 34963), "mode": 4}
        if p.get("material") is not None or p.get("tex") is not None:
            m = np.asarray(p.get("material", (1, 1, 1, 1, 1, 1)), dtype=np.float64)
            pbr = {"baseColorFactor": m[:4].tolist(), "metallicFactor": float(m[4]), "roughnessFactor": float(m[5])}
            if p.get("tex") is not None:
                buf = io.BytesIO()
                Image.fromarray(np.asarray(p["tex"], dtype=np.uint8)).save(buf, format="PNG")
                images.append({"bufferView": add_view(buf.getvalue()), "mimeType": "image/png"})
                textures.append({"source": len(images) - 1})
                pbr["baseColorTexture"] = {"index": len(textures) - 1}
            materials.append({"pbrMetallicRoughness": pbr, "doubleSided": bool(p.get("double_sided", False))})
            prim["material"] = len(materials) - 1
        gprims.append(prim)
    node = {"mesh": 0}
    if node_translation is not None:
        node["translation"] = [float(x) for x in node_translation]
    gltf = {"asset": {"version": "2.0", "generator": "rac.gltf.save_glb"}, "scene": 0, "scenes": [{"nodes": [0]}], "nodes": [node],
            "meshes": [{"primitives": gprims}], "buffers": [{"byteLength": len(blob)}], "bufferViews": views, "accessors": accessors}
    if materials:
        gltf["materials"] = materials
    if textures:
        gltf["textures"], gltf["images"] = textures, images
    js = json.dumps(gltf, separators=(",", ":")).encode("utf-8")
    js += b" " * (-len(js) % 4)
    while len(blob) % 4:
        blob.append(0)
    with open(path, "wb") as f:
        f.write(struct.pack("<4sII", b"glTF", 2, 12 + 8 + len(js) + 8 + len(blob)))
        f.write(struct.pack("<I4s", len(js), b"JSON")); f.write(js)
        f.write(struct.pack("<I4s", len(blob), b"BIN\x00")); f.write(bytes(blob))
