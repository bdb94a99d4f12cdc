// raster_shade.inl — RS-S: per-(pixel, triangle) shading (pyrender mesh.frag metallic-roughness BRDF with the
// head-light spot of renderer.py:50-57, R7/R8 in SURVEY.md §8a) and RS-D depth readback.
// Included by raster.cu (inside namespace rac).

struct TriGeom {
    int X0, Y0, X1, Y1, X2, Y2;
    float iz0, gx, gy;
};

__device__ __forceinline__ TriGeom unpack_geom(const int4 q0, const int4 q1, const int4 q2) {
    TriGeom g;
    g.X0 = q0.x; g.Y0 = q0.y; g.X1 = q0.z; g.Y1 = q0.w; g.X2 = q1.x; g.Y2 = q1.y;
    g.iz0 = __int_as_float(q1.z); g.gx = __int_as_float(q1.w); g.gy = __int_as_float(q2.x);
    return g;
}

struct ShadeCtx {
    float fx_inv, fy_inv, cx, cy;
    float iz_min, iz_max;
    float las, lao, intensity, ambient;
    float base[3];
    float metallic, alpha2, kgeo;
    const uchar4* vcolor;
    const int* faces;
    const float* uv;
    const uchar4* tex;
    int tw, th, nmips;
    const PartDev* parts;               // multi-part meshes (GLB scenes): material / texture per face
    const unsigned char* face_part;
};

__device__ __forceinline__ float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
__device__ __forceinline__ float srgb_to_linear(float c) {
    return c <= 0.04045f ? c * (1.0f / 12.92f) : __powf((c + 0.055f) * (1.0f / 1.055f), 2.4f);
}

// bilinear fetch from one mip level with GL_REPEAT wrap; returns sRGB-encoded rgb in [0,1]
__device__ __forceinline__ float3 tex_bilinear(const uchar4* __restrict__ lvl, int w, int h, float u, float v) {
    // GL texel centres at (i+0.5)/w; v is flipped by the loader so row 0 = v 0 here
    float x = u * w - 0.5f, y = v * h - 0.5f;
    const float fx0 = floorf(x), fy0 = floorf(y);
    const float ax = x - fx0, ay = y - fy0;
    int x0 = (int)fx0 % w, y0 = (int)fy0 % h;
    if (x0 < 0) x0 += w;
    if (y0 < 0) y0 += h;
    const int x1 = (x0 + 1 == w) ? 0 : x0 + 1, y1 = (y0 + 1 == h) ? 0 : y0 + 1;
    const uchar4 a = lvl[(size_t)y0 * w + x0], b = lvl[(size_t)y0 * w + x1];
    const uchar4 c = lvl[(size_t)y1 * w + x0], d = lvl[(size_t)y1 * w + x1];
    const float w00 = (1.f - ax) * (1.f - ay), w10 = ax * (1.f - ay), w01 = (1.f - ax) * ay, w11 = ax * ay;
    float3 r;
    r.x = (a.x * w00 + b.x * w10 + c.x * w01 + d.x * w11) * (1.0f / 255.0f);
    r.y = (a.y * w00 + b.y * w10 + c.y * w01 + d.y * w11) * (1.0f / 255.0f);
    r.z = (a.z * w00 + b.z * w10 + c.z * w01 + d.z * w11) * (1.0f / 255.0f);
    return r;
}

// perspective-correct barycentric weights of (v0,v1,v2) (record order) at subpixel point (Px,Py), and — GRAD — at the
// points one pixel to the right and one pixel down (GL's quad derivatives for the texture LOD).  The screen-space
// barycentrics are affine, so the neighbours follow from the edge functions' coefficients: one set of exact int64 edge
// evaluations serves all three points.
template <bool GRAD>
__device__ __forceinline__ void persp_bary(const TriGeom& g, int Px, int Py, float (&w)[3], float (&wx)[3], float (&wy)[3]) {
    const long long e0 = (long long)(g.X1 - g.X0) * (Py - g.Y0) - (long long)(g.Y1 - g.Y0) * (Px - g.X0);  // -> v2
    const long long e1 = (long long)(g.X2 - g.X1) * (Py - g.Y1) - (long long)(g.Y2 - g.Y1) * (Px - g.X1);  // -> v0
    const long long e2 = (long long)(g.X0 - g.X2) * (Py - g.Y2) - (long long)(g.Y0 - g.Y2) * (Px - g.X2);  // -> v1
    const float inv_area = 1.0f / (float)(e0 + e1 + e2);
    const float l0 = (float)e1 * inv_area, l1 = (float)e2 * inv_area, l2 = (float)e0 * inv_area;
    const float iz1 = fmaf(g.gy, (float)(g.Y1 - g.Y0), fmaf(g.gx, (float)(g.X1 - g.X0), g.iz0));
    const float iz2 = fmaf(g.gy, (float)(g.Y2 - g.Y0), fmaf(g.gx, (float)(g.X2 - g.X0), g.iz0));
    {
        const float a0 = l0 * g.iz0, a1 = l1 * iz1, a2 = l2 * iz2;
        const float s = a0 + a1 + a2;
        const float inv = (fabsf(s) > 1e-30f) ? 1.0f / s : 0.0f;
        w[0] = a0 * inv; w[1] = a1 * inv; w[2] = a2 * inv;
    }
    if (GRAD) {
        const float k = (float)kSubpix * inv_area;
        // d(e1)/dPx = -(Y2 - Y1), d(e1)/dPy = X2 - X1 (and cyclically): one pixel = kSubpix subpixels
        const float l0x = l0 - (float)(g.Y2 - g.Y1) * k, l1x = l1 - (float)(g.Y0 - g.Y2) * k, l2x = l2 - (float)(g.Y1 - g.Y0) * k;
        const float l0y = l0 + (float)(g.X2 - g.X1) * k, l1y = l1 + (float)(g.X0 - g.X2) * k, l2y = l2 + (float)(g.X1 - g.X0) * k;
        {
            const float a0 = l0x * g.iz0, a1 = l1x * iz1, a2 = l2x * iz2;
            const float s = a0 + a1 + a2;
            const float inv = (fabsf(s) > 1e-30f) ? 1.0f / s : 0.0f;
            wx[0] = a0 * inv; wx[1] = a1 * inv; wx[2] = a2 * inv;
        }
        {
            const float a0 = l0y * g.iz0, a1 = l1y * iz1, a2 = l2y * iz2;
            const float s = a0 + a1 + a2;
            const float inv = (fabsf(s) > 1e-30f) ? 1.0f / s : 0.0f;
            wy[0] = a0 * inv; wy[1] = a1 * inv; wy[2] = a2 * inv;
        }
    }
}

// COLOR_0 / texture modulation of the base colour (R7 "x COLOR_0", R8) at subpixel point (Pcx, Pcy).  Kept out of line:
// the raster warps of an SM run at unrelated program counters, so the kernel's hot instruction footprint matters.
__device__ __noinline__ void shade_base_color(const ShadeCtx& sc, const int4* __restrict__ rec, unsigned tri, int Pcx, int Pcy, float* base) {
        const TriGeom g = unpack_geom(rec[0], rec[1], rec[2]);
        float w[3], wxv[3], wyv[3];
        persp_bary<true>(g, Pcx, Pcy, w, wxv, wyv);
        const float w0 = w[0], w1 = w[1], w2 = w[2];
        // record order is (v0, v2, v1) of the mesh face (setup swaps v1 <-> v2) — except for the drawn back faces of a
        // doubleSided part, which keep the mesh order (flag: lowest mantissa bit of the record's nx, multi-part meshes only)
        const bool mesh_order = sc.parts && (rec[2].y & 1);
        const int i0 = sc.faces[3 * tri], i1 = sc.faces[3 * tri + (mesh_order ? 1 : 2)], i2 = sc.faces[3 * tri + (mesh_order ? 2 : 1)];
        // multi-part meshes: the face's own texture; COLOR_0 only where the part asks for it
        const uchar4* tex = sc.tex;
        int tw = sc.tw, th = sc.th, nmips = sc.nmips;
        bool use_vc = sc.vcolor != nullptr;
        if (sc.parts) {
            const PartDev& pt = sc.parts[sc.face_part[tri]];
            tex = pt.tex; tw = pt.tw; th = pt.th; nmips = pt.nmips;
            use_vc = use_vc && pt.has_vcolor;
        }
        if (use_vc) {
            const uchar4 c0 = sc.vcolor[i0], c1 = sc.vcolor[i1], c2 = sc.vcolor[i2];
            base[0] *= clamp01((w0 * c0.x + w1 * c1.x + w2 * c2.x) * (1.0f / 255.0f));
            base[1] *= clamp01((w0 * c0.y + w1 * c1.y + w2 * c2.y) * (1.0f / 255.0f));
            base[2] *= clamp01((w0 * c0.z + w1 * c1.z + w2 * c2.z) * (1.0f / 255.0f));
        }
        if (tex) {
            const float u = w0 * sc.uv[2 * i0] + w1 * sc.uv[2 * i1] + w2 * sc.uv[2 * i2];
            const float v = w0 * sc.uv[2 * i0 + 1] + w1 * sc.uv[2 * i1 + 1] + w2 * sc.uv[2 * i2 + 1];
            // LOD from one-pixel finite differences on the triangle's plane (GL quad derivatives)
            const float wx0 = wxv[0], wx1 = wxv[1], wx2 = wxv[2], wy0 = wyv[0], wy1 = wyv[1], wy2 = wyv[2];
            const float ux = wx0 * sc.uv[2 * i0] + wx1 * sc.uv[2 * i1] + wx2 * sc.uv[2 * i2] - u;
            const float vx = wx0 * sc.uv[2 * i0 + 1] + wx1 * sc.uv[2 * i1 + 1] + wx2 * sc.uv[2 * i2 + 1] - v;
            const float uy = wy0 * sc.uv[2 * i0] + wy1 * sc.uv[2 * i1] + wy2 * sc.uv[2 * i2] - u;
            const float vy = wy0 * sc.uv[2 * i0 + 1] + wy1 * sc.uv[2 * i1 + 1] + wy2 * sc.uv[2 * i2 + 1] - v;
            const float rx2 = ux * ux * tw * tw + vx * vx * th * th;
            const float ry2 = uy * uy * tw * tw + vy * vy * th * th;
            float lod = 0.5f * log2f(fmaxf(fmaxf(rx2, ry2), 1e-20f));
            lod = fminf(fmaxf(lod, 0.0f), (float)(nmips - 1));
            const int l0 = (int)floorf(lod);
            const int l1 = min(l0 + 1, nmips - 1);
            const float fl = lod - (float)l0;
            // walk the mip chain to level l0
            const uchar4* lvl = tex;
            int w = tw, h = th;
            for (int l = 0; l < l0; ++l) { lvl += (size_t)w * h; w = max(w >> 1, 1); h = max(h >> 1, 1); }
            float3 ta = tex_bilinear(lvl, w, h, u, v);
            if (l1 != l0) {
                const uchar4* lvl1 = lvl + (size_t)w * h;
                const float3 tb = tex_bilinear(lvl1, max(w >> 1, 1), max(h >> 1, 1), u, v);
                ta.x += (tb.x - ta.x) * fl; ta.y += (tb.y - ta.y) * fl; ta.z += (tb.z - ta.z) * fl;
            }
            base[0] *= srgb_to_linear(ta.x);
            base[1] *= srgb_to_linear(ta.y);
            base[2] *= srgb_to_linear(ta.z);
        }
}

// What the shading of one triangle needs beyond the context: its 1/z plane (relative to vertex 0) and its camera-frame
// normal.  raster_kernel keeps one per batch-local id in shared memory (the staged slot), so the untextured path never
// goes back to the 48-B record in global memory.
struct ShadeTri {
    int X0, Y0;            // vertex 0 in subpixels
    float gx, gy, iz0;
    float nx, ny, nz;
};

// RS-S: shade triangle `tri` at the centre of pixel (px,py); returns RGBA8 packed little-endian.
// `rec` (the record in global memory) is only read for vertex-coloured / textured meshes.
// PARTS: the mesh is multi-part (GLB scene) — material and texture come from the face's part.  Compiled as a separate,
// out-of-line copy (shade_parts) so that the single-material hot path keeps its register allocation.
template <bool PARTS>
__device__ __forceinline__ unsigned shade_t(const ShadeCtx& sc, const ShadeTri& t, const int4* __restrict__ rec, unsigned tri, int px, int py) {
    const float nx = t.nx, ny = t.ny, nz = t.nz;
    const int Pcx = px * kSubpix + 128, Pcy = py * kSubpix + 128;
    float iz = fmaf(t.gy, (float)(Pcy - t.Y0), fmaf(t.gx, (float)(Pcx - t.X0), t.iz0));
    iz = fminf(fmaxf(iz, sc.iz_min), sc.iz_max);
    const float z = __fdividef(1.0f, iz);   // fast-math shading: colour parity bar is 2/255 (north_star), not bit-exactness
    const float x = ((px + 0.5f) - sc.cx) * sc.fx_inv * z;
    const float y = ((py + 0.5f) - sc.cy) * sc.fy_inv * z;
    const float dist2 = x * x + y * y + z * z;
    const float inv_d = rsqrtf(dist2);
    // camera and light sit at the origin: l = v = h = -p/|p|
    const float ndl = -(nx * x + ny * y + nz * z) * inv_d;
    const float nl = fminf(fmaxf(ndl, 0.001f), 1.0f);
    const float nv = fminf(fmaxf(fabsf(ndl), 0.001f), 1.0f);
    const float nh = clamp01(ndl);
    // spot cone around the camera's forward axis (+z in the OpenCV frame)
    const float cd = clamp01(z * inv_d);
    float att = clamp01(cd * sc.las + sc.lao);
    att = att * att * (inv_d * inv_d);
    const float radiance = att * sc.intensity;

    float base[3] = {sc.base[0], sc.base[1], sc.base[2]};
    float metallic = sc.metallic, alpha2 = sc.alpha2, kk = sc.kgeo;
    if (PARTS) {
        const PartDev& pt = sc.parts[sc.face_part[tri]];
        base[0] = pt.base_color[0]; base[1] = pt.base_color[1]; base[2] = pt.base_color[2];
        metallic = pt.metallic; alpha2 = pt.alpha2; kk = pt.kgeo;
    }
    if (PARTS || sc.vcolor || sc.tex) shade_base_color(sc, rec, tri, Pcx, Pcy, base);   // out of line: cold for untextured meshes

    // metallic-roughness BRDF with v = l = h: vh = 1 -> Fresnel = f0
    const float G = __fdividef(nl, nl * (1.0f - kk) + kk) * __fdividef(nv, nv * (1.0f - kk) + kk);
    const float dd = nh * nh * (alpha2 - 1.0f) + 1.0f;
    const float D = __fdividef(alpha2, 3.14159265358979f * dd * dd);
    const float spec_common = __fdividef(G * D, 4.0f * nl * nv + 1e-4f);
    unsigned out = 0xFF000000u;
    // untextured default material: r = g = b, one gamma evaluation serves all three channels
    const bool gray = (base[0] == base[1]) && (base[1] == base[2]);
    unsigned qfirst = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        if (c > 0 && gray) { out |= qfirst << (8 * c); continue; }
        const float f0 = 0.04f + (base[c] - 0.04f) * metallic;
        const float cdiff = base[c] * 0.96f * (1.0f - metallic);
        const float diffuse = (1.0f - f0) * cdiff * (1.0f / 3.14159265358979f);
        const float col = nl * radiance * (diffuse + f0 * spec_common) + base[c] * sc.ambient;
        const float enc = clamp01(__powf(fmaxf(col, 0.0f), 1.0f / 2.2f));
        const unsigned q = (unsigned)(enc * 255.0f + 0.5f);
        if (c == 0) qfirst = q;
        out |= q << (8 * c);
    }
    return out;
}

__device__ __forceinline__ ShadeCtx make_shade_ctx(const MeshDev& mesh, const CamPose& cam, const SceneConsts& S) {
    ShadeCtx sc;
    sc.fx_inv = 1.0f / cam.fx; sc.fy_inv = 1.0f / cam.fy; sc.cx = cam.cx; sc.cy = cam.cy;
    sc.iz_min = 1.0f / S.zfar; sc.iz_max = 1.0f / S.znear;
    sc.las = S.las; sc.lao = S.lao; sc.intensity = S.intensity; sc.ambient = S.ambient;
    sc.base[0] = mesh.base_color[0]; sc.base[1] = mesh.base_color[1]; sc.base[2] = mesh.base_color[2];
    sc.metallic = mesh.metallic;
    const float rough = fminf(fmaxf(mesh.roughness, 0.04f), 1.0f);
    const float alpha = rough * rough;
    sc.alpha2 = alpha * alpha;
    sc.kgeo = (rough + 1.0f) * (rough + 1.0f) * 0.125f;
    sc.vcolor = mesh.vcolor; sc.faces = mesh.faces; sc.uv = mesh.uv; sc.tex = mesh.tex;
    sc.tw = mesh.tw; sc.th = mesh.th; sc.nmips = mesh.nmips;
    sc.parts = mesh.parts; sc.face_part = mesh.face_part;
    return sc;
}

// RS-D
__device__ __forceinline__ float depth_readback(unsigned d24, const SceneConsts& S) {
    const float d = __fdiv_rn((float)d24, 16777215.0f);
    const float zn = __fsub_rn(__fmul_rn(2.0f, d), 1.0f);
    return __fdiv_rn(S.rb_num, __fsub_rn(S.rb_sum, __fmul_rn(zn, S.rb_diff)));
}
