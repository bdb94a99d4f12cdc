"""Mesh loading, procedural meshes and the device mesh cache.

Replaces the `tm.load(path)` + `pyrender.Mesh.from_trimesh(smooth=False)` pair the reference
runs for EVERY rendered image (renderer.py:28-48, called from renderer.py:99) and the vertex
source of data_loaders.py:19-32: a mesh file is parsed once, uploaded once
(rac_mesh_create) and afterwards referred to by an integer handle.
"""
import ctypes
import os
import struct
import threading

import numpy as np

from . import _lib

_PLY_TYPES = {
    "char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
    "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
    "float": "f4", "float32": "f4", "double": "f8", "float64": "f8",
}


VERTEX_COLOR_MATERIAL = (1.0, 1.0, 1.0, 1.0, 0.2, 0.8)     # pyrender: material of a vertex-/face-coloured trimesh without a material


class Mesh:
    """Indexed triangle mesh on the host.  verts (V,3) f32, faces (T,3) i32, optional
    uv (V,2) f32, vcolor (V,4) u8, tex (th,tw,3) u8 with row 0 at v = 0 (GL convention)."""

    def __init__(self, verts, faces, uv=None, vcolor=None, tex=None, material=None):
        self.verts = np.ascontiguousarray(verts, dtype=np.float32).reshape(-1, 3)
        self.faces = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        self.uv = None if uv is None else np.ascontiguousarray(uv, dtype=np.float32).reshape(-1, 2)
        self.vcolor = None if vcolor is None else np.ascontiguousarray(vcolor, dtype=np.uint8).reshape(-1, 4)
        self.tex = None if tex is None else np.ascontiguousarray(tex, dtype=np.uint8)
        # (base_color[4], metallic, roughness); None = pyrender's default material (R7)
        self.material = material
        # multi-part meshes (GLB scenes, rac/gltf.py): part index per face + [dict(material, tex, use_vcolor)] per part
        self.face_part = None
        self.parts = None
        assert self.faces.size == 0 or (self.faces.min() >= 0 and self.faces.max() < len(self.verts))

    @property
    def vertices(self):   # trimesh-compatible name (data_loaders.py:21)
        return self.verts

    def material_array(self):
        """(base_color[4], metallic, roughness) of a single-part mesh.  Without an explicit material the choice follows
        pyrender's Mesh.from_trimesh -> _get_trimesh_props (renderer.py:41): a mesh that carries vertex colours
        (trimesh visual kind 'vertex', e.g. a coloured PLY) gets baseColorFactor (1, 1, 1, 1) — the vertex colour IS the
        base colour — and only a mesh with no colours at all gets the 0.3 grey.  Both use metallic 0.2, roughness 0.8."""
        if self.material is None:
            if self.vcolor is not None and self.tex is None:
                return np.array(VERTEX_COLOR_MATERIAL, dtype=np.float32)
            return np.array([0.3, 0.3, 0.3, 1.0, 0.2, 0.8], dtype=np.float32)
        return np.asarray(self.material, dtype=np.float32).reshape(6)


# ---- PLY --------------------------------------------------------------------------------------
def _parse_header(f):
    line = f.readline().strip()
    if line != b"ply":
        raise ValueError("not a PLY file")
    fmt = None
    elements = []   # [name, count, [(prop_name, dtype) | (prop_name, (count_t, item_t))]]
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header not terminated")
        tok = line.strip().split()
        if not tok or tok[0] in (b"comment", b"obj_info"):
            continue
        if tok[0] == b"format":
            fmt = tok[1].decode()
        elif tok[0] == b"element":
            elements.append([tok[1].decode(), int(tok[2]), []])
        elif tok[0] == b"property":
            if tok[1] == b"list":
                elements[-1][2].append((tok[4].decode(), (_PLY_TYPES[tok[2].decode()], _PLY_TYPES[tok[3].decode()])))
            else:
                elements[-1][2].append((tok[2].decode(), _PLY_TYPES[tok[1].decode()]))
        elif tok[0] == b"end_header":
            break
    return fmt, elements


def load_ply(path):
    """ASCII / binary PLY with optional vertex colours and uv.  Polygons are fan-triangulated."""
    with open(path, "rb") as f:
        fmt, elements = _parse_header(f)
        data = {}
        if fmt == "ascii":
            tokens = f.read().split()
            pos = 0
            for name, count, props in elements:
                has_list = any(isinstance(t, tuple) for _, t in props)
                if not has_list:
                    n = count * len(props)
                    arr = np.array(tokens[pos:pos + n], dtype=np.float64).reshape(count, len(props))
                    pos += n
                    data[name] = {p: arr[:, i] for i, (p, _) in enumerate(props)}
                else:
                    rows = []
                    for _ in range(count):
                        row = {}
                        for p, t in props:
                            if isinstance(t, tuple):
                                k = int(tokens[pos]); pos += 1
                                row[p] = [int(float(x)) for x in tokens[pos:pos + k]]; pos += k
                            else:
                                row[p] = float(tokens[pos]); pos += 1
                        rows.append(row)
                    data[name] = rows
        else:
            end = "<" if fmt == "binary_little_endian" else ">"
            for name, count, props in elements:
                has_list = any(isinstance(t, tuple) for _, t in props)
                if not has_list:
                    dt = np.dtype([(p, end + t) for p, t in props])
                    arr = np.frombuffer(f.read(dt.itemsize * count), dtype=dt, count=count)
                    data[name] = {p: arr[p] for p, _ in props}
                else:
                    rows = []
                    for _ in range(count):
                        row = {}
                        for p, t in props:
                            if isinstance(t, tuple):
                                ct, it = np.dtype(end + t[0]), np.dtype(end + t[1])
                                k = int(np.frombuffer(f.read(ct.itemsize), dtype=ct)[0])
                                row[p] = np.frombuffer(f.read(it.itemsize * k), dtype=it).astype(np.int64).tolist()
                            else:
                                d = np.dtype(end + t)
                                row[p] = float(np.frombuffer(f.read(d.itemsize), dtype=d)[0])
                        rows.append(row)
                    data[name] = rows
    v = data["vertex"]
    verts = np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32)
    vcolor = None
    if all(k in v for k in ("red", "green", "blue")):
        a = v["alpha"] if "alpha" in v else np.full(len(verts), 255)
        vcolor = np.stack([v["red"], v["green"], v["blue"], a], axis=1).astype(np.uint8)
    uv = None
    for ku, kv in (("s", "t"), ("u", "v"), ("texture_u", "texture_v")):
        if ku in v and kv in v:
            uv = np.stack([v[ku], v[kv]], axis=1).astype(np.float32)
            break
    faces = []
    for row in data.get("face", []):
        idx = row.get("vertex_indices", row.get("vertex_index"))
        for k in range(1, len(idx) - 1):
            faces.append((idx[0], idx[k], idx[k + 1]))
    return Mesh(verts, np.asarray(faces, dtype=np.int32).reshape(-1, 3), uv=uv, vcolor=vcolor)


def save_ply(path, mesh, binary=True):
    n, t = len(mesh.verts), len(mesh.faces)
    props = ["property float x", "property float y", "property float z"]
    if mesh.vcolor is not None:
        props += ["property uchar red", "property uchar green", "property uchar blue", "property uchar alpha"]
    if mesh.uv is not None:
        props += ["property float s", "property float t"]
    header = ["ply", "format %s 1.0" % ("binary_little_endian" if binary else "ascii"), "element vertex %d" % n] + props + [
        "element face %d" % t, "property list uchar int vertex_indices", "end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode())
        for i in range(n):
            if binary:
                f.write(struct.pack("<3f", *mesh.verts[i]))
                if mesh.vcolor is not None:
                    f.write(struct.pack("<4B", *mesh.vcolor[i]))
                if mesh.uv is not None:
                    f.write(struct.pack("<2f", *mesh.uv[i]))
            else:
                row = ["%.9g" % x for x in mesh.verts[i]]
                if mesh.vcolor is not None:
                    row += [str(int(x)) for x in mesh.vcolor[i]]
                if mesh.uv is not None:
                    row += ["%.9g" % x for x in mesh.uv[i]]
                f.write((" ".join(row) + "\n").encode())
        for i in range(t):
            if binary:
                f.write(struct.pack("<B3i", 3, *mesh.faces[i]))
            else:
                f.write(("3 %d %d %d\n" % tuple(mesh.faces[i])).encode())


def load_mesh(path):
    ext = os.path.splitext(path)[1].lower()
    if ext == ".ply":
        return load_ply(path)
    if ext == ".npz":   # this repo's own container for textured synthetic meshes
        z = np.load(path)
        get = lambda k: z[k] if k in z.files else None
        return Mesh(z["verts"], z["faces"], uv=get("uv"), vcolor=get("vcolor"), tex=get("tex"), material=get("material"))
    if ext == ".glb":   # the reference's "texturized" datasets (renderer.py:36-46)
        from .gltf import load_glb
        return load_glb(path)
    raise ValueError(f"unsupported mesh format {ext!r} (PLY, GLB and NPZ are supported)")


# ---- procedural meshes (synthetic inputs: the reference's datasets are unavailable) ----
def icosphere(level=3, radius=0.5):
    """Outward-wound icosphere: 20*4^level triangles."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
         (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
         (8, 6, 7), (9, 8, 1)]
    verts = [np.array(p, dtype=np.float64) / np.linalg.norm(p) for p in v]
    faces = list(f)
    for _ in range(level):
        cache = {}

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = verts[a] + verts[b]
                verts.append(m / np.linalg.norm(m))
                cache[key] = len(verts) - 1
            return cache[key]

        nf = []
        for a, b, c in faces:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        faces = nf
    return Mesh(np.array(verts) * radius, np.array(faces, dtype=np.int32))


def bumpy_sphere(level=5, radius=0.45, seed=0, amp=0.18, decimate_to=None):
    """Seeded radial-noise icosphere ("ModelNet10-scale" stand-in): smooth low-frequency lobes so the
    silhouette is non-convex and faces occlude each other.  decimate_to drops random faces (holes are
    reference behaviour for badly wound ModelNet meshes too)."""
    m = icosphere(level, 1.0)
    rng = np.random.default_rng(seed)
    d = m.verts.astype(np.float64)
    r = np.ones(len(d))
    for _ in range(6):
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        freq = rng.uniform(1.5, 4.0)
        r += amp * rng.uniform(0.3, 1.0) * np.sin(freq * (d @ axis) * np.pi + rng.uniform(0, 2 * np.pi))
    verts = d * (radius * r / r.max())[:, None]
    faces = m.faces
    if decimate_to is not None and decimate_to < len(faces):
        keep = np.sort(rng.choice(len(faces), size=decimate_to, replace=False))
        faces = faces[keep]
    return Mesh(verts, faces)


def box(sx=0.6, sy=0.4, sz=0.3, subdiv=1):
    """Axis-aligned box, outward winding, each face split into subdiv x subdiv x 2 triangles."""
    verts, faces = [], []
    h = np.array([sx, sy, sz]) / 2.0
    for axis in range(3):
        for sign in (-1, 1):
            u, v = (axis + 1) % 3, (axis + 2) % 3
            base = len(verts)
            for i in range(subdiv + 1):
                for j in range(subdiv + 1):
                    p = np.zeros(3)
                    p[axis] = sign * h[axis]
                    p[u] = -h[u] + 2 * h[u] * i / subdiv
                    p[v] = -h[v] + 2 * h[v] * j / subdiv
                    verts.append(p)
            for i in range(subdiv):
                for j in range(subdiv):
                    a = base + i * (subdiv + 1) + j
                    b, c, d = a + subdiv + 1, a + subdiv + 2, a + 1
                    if sign > 0:
                        faces += [(a, b, c), (a, c, d)]
                    else:
                        faces += [(a, c, b), (a, d, c)]
    return Mesh(np.array(verts), np.array(faces, dtype=np.int32))


def merge(meshes):
    verts, faces, off = [], [], 0
    for m in meshes:
        verts.append(m.verts)
        faces.append(m.faces + off)
        off += len(m.verts)
    return Mesh(np.concatenate(verts), np.concatenate(faces))


def table(subdiv=8):
    """Box-union 'table': a top and four legs (interpenetrating parts, like ModelNet furniture)."""
    parts = [box(0.9, 0.6, 0.06, subdiv)]
    parts[0].verts[:, 2] += 0.25
    for sx in (-0.38, 0.38):
        for sy in (-0.24, 0.24):
            leg = box(0.07, 0.07, 0.5, max(1, subdiv // 2))
            leg.verts += np.array([sx, sy, 0.0], dtype=np.float32)
            parts.append(leg)
    return merge(parts)


def with_texture(mesh, size=256, seed=0):
    """Spherical-uv + seeded checker/noise RGB8 texture (texturised-ModelNet stand-in, R8).
    Textured pyrender meshes get baseColorFactor 1, metallic 0.2?  No: trimesh SimpleMaterial ->
    MetallicRoughnessMaterial(baseColorFactor=diffuse, roughness from Ns); we use (1,1,1,1), 0.2, 0.8."""
    rng = np.random.default_rng(seed)
    d = mesh.verts / np.maximum(np.linalg.norm(mesh.verts, axis=1, keepdims=True), 1e-9)
    uv = np.stack([np.arctan2(d[:, 1], d[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(d[:, 2], -1, 1)) / np.pi], axis=1)
    yy, xx = np.mgrid[0:size, 0:size]
    checker = (((xx // (size // 8)) + (yy // (size // 8))) % 2).astype(np.float64)
    tex = np.empty((size, size, 3), dtype=np.float64)
    tex[..., 0] = 80 + 120 * checker
    tex[..., 1] = 60 + 150 * (xx / size)
    tex[..., 2] = 200 - 120 * checker * (yy / size)
    tex += rng.uniform(-20, 20, size=tex.shape)
    return Mesh(mesh.verts, mesh.faces, uv=uv, tex=np.clip(tex, 0, 255).astype(np.uint8),
                material=np.array([1, 1, 1, 1, 0.2, 0.8], dtype=np.float32))


# ---- device cache ------------------------------------------------------------------------------
class DeviceMesh:
    """A mesh uploaded to the GPU (rac_mesh_create).  `handle` indexes the library's device table."""

    def __init__(self, mesh):
        _lib.require_cuda()
        L = _lib.lib()
        self.mesh = mesh
        self.V, self.T = len(mesh.verts), len(mesh.faces)
        h = ctypes.c_int(-1)
        vp = lambda a: a.ctypes.data_as(ctypes.c_void_p) if a is not None else None
        if getattr(mesh, "parts", None):
            n = len(mesh.parts)
            arr = (_lib.RacPart * n)()
            self._keep = []                     # texture arrays must outlive the call
            for k, p in enumerate(mesh.parts):
                m = np.asarray(p["material"], dtype=np.float32).reshape(6)
                for i in range(4):
                    arr[k].material.base_color[i] = float(m[i])
                arr[k].material.metallic, arr[k].material.roughness = float(m[4]), float(m[5])
                arr[k].use_vertex_color = int(bool(p.get("use_vcolor", False)))
                arr[k].double_sided = int(bool(p.get("double_sided", False)))
                tex = p.get("tex")
                if tex is not None and mesh.uv is not None:
                    tex = np.ascontiguousarray(tex, dtype=np.uint8)
                    self._keep.append(tex)
                    arr[k].tex_rgb8 = tex.ctypes.data_as(ctypes.c_void_p)
                    arr[k].tex_h, arr[k].tex_w = tex.shape[:2]
            fp = np.ascontiguousarray(mesh.face_part, dtype=np.uint8)
            rc = L.rac_mesh_create_parts(vp(mesh.verts), self.V, vp(mesh.faces), self.T, vp(mesh.uv), vp(mesh.vcolor),
                                         vp(fp), n, arr, ctypes.byref(h))
        else:
            mat = _lib.RacMaterial()
            arr = mesh.material_array()
            for i in range(4):
                mat.base_color[i] = float(arr[i])
            mat.metallic, mat.roughness = float(arr[4]), float(arr[5])
            tw = th = 0
            tex_p = None
            if mesh.tex is not None and mesh.uv is not None:
                th, tw = mesh.tex.shape[:2]
                tex_p = mesh.tex.ctypes.data_as(ctypes.c_void_p)
            rc = L.rac_mesh_create(vp(mesh.verts), self.V, vp(mesh.faces), self.T, vp(mesh.uv), tex_p, tw, th,
                                   vp(mesh.vcolor), ctypes.byref(mat), ctypes.byref(h))
        _lib.check(rc)
        self.handle = h.value

    def destroy(self):
        if self.handle >= 0:
            _lib.check(_lib.lib().rac_mesh_destroy(self.handle))
            self.handle = -1


# Path / object -> uploaded mesh, least recently used first.  The reference re-parses the file for every image and keeps
# nothing; training draws from thousands of distinct ModelNet10 files (roughly 4.7 k across train + validation), more
# than the library's device table holds (kMaxMeshes = 4096, rac_common.cuh), so the cache is BOUNDED: beyond
# MAX_DEVICE_MESHES entries the least recently used meshes are destroyed and simply uploaded again if they are ever
# drawn again.  Rules of eviction:
#   * only meshes nobody else holds: a DeviceMesh a caller still references (rac.RenderCompareStep.meshes, a captured
#     graph's owner — mesh ids are baked into graphs by value, and a freed slot is reused by the next upload) is
#     skipped, so a handle stays valid for as long as its DeviceMesh object is alive outside the cache;
#   * never while the current stream is capturing (rac_mesh_destroy synchronises the device — a batch in flight may still
#     read the mesh — which is illegal under capture): the cache overshoots instead, the device table has 4x the room;
#   * _EVICT_BATCH victims at a time, so the device-wide synchronise is paid once per batch of evictions, not per mesh
#     (the second and later destroys of a batch find the device idle).
import collections
import sys

MAX_DEVICE_MESHES = 1024
MAX_HOST_MESHES = 4096
_EVICT_BATCH = 64
_cache = collections.OrderedDict()
_cache_lock = threading.Lock()
_host_cache = collections.OrderedDict()


def _capturing():
    try:
        import torch
        return torch.cuda.is_available() and torch.cuda.is_current_stream_capturing()
    except Exception:
        return False


def _unheld_refcount():
    """sys.getrefcount of a value that only a dict holds, measured the way _evict_lru asks (interpreter-version proof)."""
    d = {0: object()}
    return sys.getrefcount(d[0])


_UNHELD = _unheld_refcount()


def _evict_lru():
    victims = []
    for k in _cache:                              # keys, least recently used first (no loop variable holds a value)
        if sys.getrefcount(_cache[k]) <= _UNHELD: # only the cache references it
            victims.append(k)
            if len(victims) >= _EVICT_BATCH:
                break
    for k in victims:
        _cache.pop(k).destroy()


def _cache_get(key, make, owner=None):
    """LRU lookup / insert under _cache_lock (the caller holds it).  owner: the host Mesh of an ("obj", id) key; the
    DeviceMesh keeps it alive (dm.mesh), so its id() cannot be recycled while the entry exists — checked all the same."""
    dm = _cache.get(key)
    if dm is not None and (owner is None or dm.mesh is owner):
        _cache.move_to_end(key)
        return dm
    if len(_cache) >= MAX_DEVICE_MESHES and not _capturing():
        _evict_lru()
    dm = _cache[key] = make()
    return dm


def host_mesh(path):
    """Parsed-once host mesh by path."""
    key = os.path.abspath(path)
    with _cache_lock:
        hm = _host_cache.get(key)
        if hm is None:
            while len(_host_cache) >= MAX_HOST_MESHES:
                _host_cache.popitem(last=False)
            hm = _host_cache[key] = load_mesh(key)
        else:
            _host_cache.move_to_end(key)
        return hm


def device_mesh(path_or_mesh):
    """Uploaded-once device mesh by path (or by Mesh object identity)."""
    if isinstance(path_or_mesh, DeviceMesh):
        return path_or_mesh
    if isinstance(path_or_mesh, Mesh):
        with _cache_lock:
            return _cache_get(("obj", id(path_or_mesh)), lambda: DeviceMesh(path_or_mesh), owner=path_or_mesh)
    key = os.path.abspath(path_or_mesh)
    hm = host_mesh(key)
    with _cache_lock:
        return _cache_get(key, lambda: DeviceMesh(hm))


def clear_cache():
    with _cache_lock:
        for dm in _cache.values():
            dm.destroy()
        _cache.clear()
        _host_cache.clear()
