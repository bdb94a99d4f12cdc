/*
 * raster_ref.c — CPU ORACLE for steps 1-2 of the render-and-compare hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (the package
 * end-to-end-rgb-pose-estimation-baseline_b200/) may import, link or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * PARITY UNPINNED.  The reference renders through pyrender + OpenGL/EGL
 * (reference renderer.py:3,9,77-82), which is neither under /root/reference nor installable
 * nor runnable here (no graphics stack), and the reference ships no test, golden image or
 * fixture for this path.  This file restates, in plain C, the pipeline that
 * renderer.py:85-105 (render_scene) drives inside pyrender 0.1.45, as recalled in SURVEY.md
 * §8a R1-R9 (camera.py IntrinsicsCamera.get_projection_matrix, renderer.py
 * _configure_main_framebuffer / _read_main_framebuffer, shaders/mesh.vert|frag), plus the
 * glue around it: data_loaders.py:100-118 (render_batch_sequential) and
 * data_loaders.py:69-82,158-177 (normalize_depth, normalize_img, prepare_model_input).
 *
 * What each function follows:
 *   project()            R1 (renderer.py:96 CV->GL flip, folded: we stay in the CV frame) and
 *                        R2 (IntrinsicsCamera projection: u = fx x/z + cx, v = fy y/z + cy,
 *                        pixel centres at +0.5); GL viewport transform snaps to 8 subpixel bits.
 *   raster loop          R4 (single-sided material -> GL_CULL_FACE back, GL_LESS, 24-bit depth),
 *                        R5 (4x MSAA renderbuffers, standard sample pattern, top-left rule).
 *   shade()              R7 (mesh.frag metallic-roughness branch, SpotLight at the camera pose:
 *                        renderer.py:50-57), R8 (baseColorTexture, sRGB decode, trilinear+repeat).
 *   resolve              R5 (glBlitFramebuffer: colour = mean of 4 samples, depth = sample 0).
 *   depth read-back      R6 (pyrender _read_main_framebuffer linearisation, f32 numpy ops).
 *   oracle_render_batch  renderer.py:103-104 + data_loaders.py:112-114 (max-normalise, f32).
 *   oracle_prepare_model_input   data_loaders.py:69-82,158-177.
 *
 * Numerics: the coverage/depth part is specified op-by-op (fmaf where the spec says fma,
 * separate IEEE ops elsewhere; build with -ffp-contract=off) so that a second implementation
 * following the same spec reproduces depth and coverage bit for bit.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define EXPORT __attribute__((visibility("default")))

typedef struct {
    const float* verts;      /* V*3 */
    const int32_t* faces;    /* T*3 */
    int V, T;
    const float* uv;         /* V*2 or NULL */
    const uint8_t* vcolor;   /* V*4 or NULL */
    const uint8_t* tex;      /* th*tw*3 RGB8 (row 0 = v 0) or NULL */
    int tw, th;
    float base_color[4];
    float metallic, roughness;
    /* multi-part meshes (a .glb scene, renderer.py:36-42: every geometry is its own pyrender mesh with the file's
     * material, all in one scene): per-face part index; per part a material (6 floats), a texture and whether COLOR_0
     * applies (pyrender: only primitives without a material) */
    const uint8_t* face_part;     /* T or NULL */
    int nparts;
    const float* part_material;   /* nparts*6 */
    const int* part_use_vcolor;   /* nparts */
    const int* part_double_sided; /* nparts: glTF doubleSided — back faces are drawn (normal not flipped) */
    const int* part_tw;           /* nparts */
    const int* part_th;
    const struct Mips_s* part_mips;
} OMesh;

typedef struct {
    float ambient, intensity, inner_cone, outer_cone, znear, zfar;
} OScene;

typedef struct {
    int X[3], Y[3];
    float iz[3];
} Proj;

/* mip chain for R8: box filter on the encoded 8-bit values (GL_RGBA8 texture, glGenerateMipmap) */
typedef struct Mips_s {
    int n;
    int w[16], h[16];
    uint8_t* lvl[16];   /* RGB8 */
} Mips;

static void mips_build(Mips* m, const uint8_t* tex, int tw, int th) {
    m->n = 0;
    if (!tex) return;
    int w = tw, h = th;
    uint8_t* cur = (uint8_t*)malloc((size_t)w * h * 3);
    memcpy(cur, tex, (size_t)w * h * 3);
    for (;;) {
        m->w[m->n] = w; m->h[m->n] = h; m->lvl[m->n] = cur; m->n++;
        if ((w == 1 && h == 1) || m->n == 16) break;
        int nw = w > 1 ? w / 2 : 1, nh = h > 1 ? h / 2 : 1;
        uint8_t* nx = (uint8_t*)malloc((size_t)nw * nh * 3);
        for (int y = 0; y < nh; ++y)
            for (int x = 0; x < nw; ++x) {
                int xa = 2 * x < w ? 2 * x : w - 1, xb = 2 * x + 1 < w ? 2 * x + 1 : w - 1;
                int ya = 2 * y < h ? 2 * y : h - 1, yb = 2 * y + 1 < h ? 2 * y + 1 : h - 1;
                for (int c = 0; c < 3; ++c) {
                    int s = cur[((size_t)ya * w + xa) * 3 + c] + cur[((size_t)ya * w + xb) * 3 + c] +
                            cur[((size_t)yb * w + xa) * 3 + c] + cur[((size_t)yb * w + xb) * 3 + c];
                    nx[((size_t)y * nw + x) * 3 + c] = (uint8_t)((s + 2) >> 2);
                }
            }
        cur = nx; w = nw; h = nh;
    }
}
static void mips_free(Mips* m) {
    for (int i = 0; i < m->n; ++i) free(m->lvl[i]);
    m->n = 0;
}

/* R1+R2: camera-frame point, perspective divide, pixel coordinates, 1/256-pixel snap.
 * Returns 0 when the vertex kills its triangle (behind the near plane / outside the guard band). */
static int project(const float* T, const float* K, const float* p, float znear, int* X, int* Y, float* iz) {
    float xc = fmaf(T[2], p[2], fmaf(T[1], p[1], fmaf(T[0], p[0], T[3])));
    float yc = fmaf(T[6], p[2], fmaf(T[5], p[1], fmaf(T[4], p[0], T[7])));
    float zc = fmaf(T[10], p[2], fmaf(T[9], p[1], fmaf(T[8], p[0], T[11])));
    if (!(zc >= znear * (1.0f / 64.0f))) return 0;   /* RS-V: projectable down to znear / 64; the near plane clips per sample (d < 0) */
    float inv = 1.0f / zc;
    float xn = xc * inv, yn = yc * inv;
    float u = fmaf(K[0], xn, K[2]);
    float v = fmaf(K[4], yn, K[5]);
    if (!(fabsf(u) <= 16384.0f) || !(fabsf(v) <= 16384.0f)) return 0;
    *X = (int)lrintf(u * 256.0f);
    *Y = (int)lrintf(v * 256.0f);
    *iz = inv;
    return 1;
}

static int cdiv256(int a) { return (a + 255) >> 8; }   /* ceil(a/256), arithmetic shift */
static int fdiv256(int a) { return a >> 8; }           /* floor(a/256) */

/* MSAA 4x standard sample positions in 1/256 pixel */
static const int SMP_X[4] = {96, 224, 32, 160};
static const int SMP_Y[4] = {32, 96, 160, 224};

typedef struct {
    /* oriented so that the interior is on the >= 0 side of every edge */
    int X[3], Y[3];
    float iz0, gx, gy;
    int mesh_order;   /* 1: vertices kept in mesh order (drawn back face of a doubleSided part), 0: v1 <-> v2 swapped */
} Setup;

/* returns 1 if the triangle is front-facing, non-degenerate and inside the depth/guard limits */
static int tri_setup(const OMesh* m, int t, const float* T, const float* K, float znear, Setup* s) {
    Proj q;
    for (int k = 0; k < 3; ++k)
        if (!project(T, K, m->verts + 3 * (size_t)m->faces[3 * t + k], znear, &q.X[k], &q.Y[k], &q.iz[k])) return 0;
    int64_t area2 = (int64_t)(q.X[1] - q.X[0]) * (q.Y[2] - q.Y[0]) - (int64_t)(q.X[2] - q.X[0]) * (q.Y[1] - q.Y[0]);
    int back = 0;
    if (area2 > 0 && m->face_part && m->part_double_sided) back = m->part_double_sided[m->face_part[t]] != 0;
    if (area2 >= 0 && !back) return 0;   /* R4: front = CCW in GL window space = negative area with v down */
    if (area2 == 0) return 0;
    const int a = back ? 1 : 2, b = back ? 2 : 1;      /* interior on the >= 0 side: swap v1 <-> v2 for front faces */
    s->mesh_order = back;
    s->X[0] = q.X[0]; s->Y[0] = q.Y[0];
    s->X[1] = q.X[a]; s->Y[1] = q.Y[a];
    s->X[2] = q.X[b]; s->Y[2] = q.Y[b];
    float iz1 = q.iz[a], iz2 = q.iz[b];
    s->iz0 = q.iz[0];
    float area = (float)(back ? area2 : -area2);
    float d1 = iz1 - s->iz0, d2 = iz2 - s->iz0;
    float ex1 = (float)(s->X[1] - s->X[0]), ey1 = (float)(s->Y[1] - s->Y[0]);
    float ex2 = (float)(s->X[2] - s->X[0]), ey2 = (float)(s->Y[2] - s->Y[0]);
    float t1 = d2 * ey1, t2 = d1 * ex2;
    s->gx = fmaf(d1, ey2, -t1) / area;
    s->gy = fmaf(d2, ex1, -t2) / area;
    return 1;
}

static float clamp01(float x) { return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x); }
static float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
static float srgb_decode(float c) { return c <= 0.04045f ? c / 12.92f : powf((c + 0.055f) / 1.055f, 2.4f); }

static void pbary(const Setup* s, int Px, int Py, float w[3]) {
    int64_t e0 = (int64_t)(s->X[1] - s->X[0]) * (Py - s->Y[0]) - (int64_t)(s->Y[1] - s->Y[0]) * (Px - s->X[0]);
    int64_t e1 = (int64_t)(s->X[2] - s->X[1]) * (Py - s->Y[1]) - (int64_t)(s->Y[2] - s->Y[1]) * (Px - s->X[1]);
    int64_t e2 = (int64_t)(s->X[0] - s->X[2]) * (Py - s->Y[2]) - (int64_t)(s->Y[0] - s->Y[2]) * (Px - s->X[2]);
    float ia = 1.0f / (float)(e0 + e1 + e2);
    float l0 = (float)e1 * ia, l1 = (float)e2 * ia, l2 = (float)e0 * ia;
    float iz1 = fmaf(s->gy, (float)(s->Y[1] - s->Y[0]), fmaf(s->gx, (float)(s->X[1] - s->X[0]), s->iz0));
    float iz2 = fmaf(s->gy, (float)(s->Y[2] - s->Y[0]), fmaf(s->gx, (float)(s->X[2] - s->X[0]), s->iz0));
    float a0 = l0 * s->iz0, a1 = l1 * iz1, a2 = l2 * iz2;
    float sum = a0 + a1 + a2;
    float inv = fabsf(sum) > 1e-30f ? 1.0f / sum : 0.0f;
    w[0] = a0 * inv; w[1] = a1 * inv; w[2] = a2 * inv;
}

static void tex_fetch(const Mips* mp, int l, float u, float v, float out[3]) {
    int w = mp->w[l], h = mp->h[l];
    const uint8_t* d = mp->lvl[l];
    float x = u * w - 0.5f, y = v * h - 0.5f;
    float fx0 = floorf(x), fy0 = floorf(y);
    float ax = x - fx0, ay = y - fy0;
    int x0 = (int)fx0 % w, y0 = (int)fy0 % h;
    if (x0 < 0) x0 += w;
    if (y0 < 0) y0 += h;
    int x1 = x0 + 1 == w ? 0 : x0 + 1, y1 = y0 + 1 == h ? 0 : y0 + 1;
    float w00 = (1 - ax) * (1 - ay), w10 = ax * (1 - ay), w01 = (1 - ax) * ay, w11 = ax * ay;
    for (int c = 0; c < 3; ++c)
        out[c] = (d[((size_t)y0 * w + x0) * 3 + c] * w00 + d[((size_t)y0 * w + x1) * 3 + c] * w10 +
                  d[((size_t)y1 * w + x0) * 3 + c] * w01 + d[((size_t)y1 * w + x1) * 3 + c] * w11) / 255.0f;
}

/* R7/R8: one fragment-shader invocation at the centre of pixel (px,py) for triangle t */
static void shade(const OMesh* m, const Mips* mp, const OScene* sc, const float* T, const float* K, int t,
                  const Setup* s, int px, int py, uint8_t rgba[4]) {
    const float* a = m->verts + 3 * (size_t)m->faces[3 * t];
    const float* b = m->verts + 3 * (size_t)m->faces[3 * t + 1];
    const float* c = m->verts + 3 * (size_t)m->faces[3 * t + 2];
    /* R3: flat face normal, object frame (double, like trimesh), rotated into the camera frame */
    double e1[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]};
    double e2[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
    double nd[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
    double len = sqrt(nd[0] * nd[0] + nd[1] * nd[1] + nd[2] * nd[2]);
    float no[3] = {0, 0, 0};
    if (len > 0.0) { no[0] = (float)(nd[0] / len); no[1] = (float)(nd[1] / len); no[2] = (float)(nd[2] / len); }
    float n[3] = {T[0] * no[0] + T[1] * no[1] + T[2] * no[2], T[4] * no[0] + T[5] * no[1] + T[6] * no[2],
                  T[8] * no[0] + T[9] * no[1] + T[10] * no[2]};

    int Pcx = px * 256 + 128, Pcy = py * 256 + 128;
    float iz = fmaf(s->gy, (float)(Pcy - s->Y[0]), fmaf(s->gx, (float)(Pcx - s->X[0]), s->iz0));
    iz = clampf(iz, 1.0f / sc->zfar, 1.0f / sc->znear);
    float z = 1.0f / iz;
    float x = ((px + 0.5f) - K[2]) / K[0] * z;
    float y = ((py + 0.5f) - K[5]) / K[4] * z;
    float dist2 = x * x + y * y + z * z;
    float dist = sqrtf(dist2);
    /* camera position == light position == origin of the camera frame */
    float l[3] = {-x / dist, -y / dist, -z / dist};
    float ndl = n[0] * l[0] + n[1] * l[1] + n[2] * l[2];
    float nl = clampf(ndl, 0.001f, 1.0f);
    float nv = clampf(fabsf(ndl), 0.001f, 1.0f);
    float nh = clamp01(ndl);   /* h = normalize(l + v) = l */
    /* spot attenuation: light axis = camera forward = +z in the CV frame */
    float las = 1.0f / fmaxf(0.001f, cosf(sc->inner_cone) - cosf(sc->outer_cone));
    float lao = -cosf(sc->outer_cone) * las;
    float cd = clamp01(-l[2]);
    float att = clamp01(cd * las + lao);
    att = att * att / dist2;
    float radiance = att * sc->intensity;

    const float* bc = m->base_color;
    float metallic = m->metallic, roughness = m->roughness;
    int use_vc = m->vcolor != NULL, tw = m->tw, th = m->th;
    if (m->face_part) {
        const int part = m->face_part[t];
        bc = m->part_material + 6 * part;
        metallic = bc[4]; roughness = bc[5];
        mp = &m->part_mips[part];
        tw = m->part_tw[part]; th = m->part_th[part];
        use_vc = use_vc && m->part_use_vcolor[part];
    }
    float base[3] = {bc[0], bc[1], bc[2]};
    if (use_vc || mp->n) {
        float w[3];
        pbary(s, Pcx, Pcy, w);
        /* Setup order is (v0, v2, v1) of the mesh face, (v0, v1, v2) for a drawn back face */
        int id[3] = {m->faces[3 * t], m->faces[3 * t + (s->mesh_order ? 1 : 2)], m->faces[3 * t + (s->mesh_order ? 2 : 1)]};
        if (use_vc)
            for (int k = 0; k < 3; ++k) {
                float v = 0.0f;
                for (int j = 0; j < 3; ++j) v += w[j] * (float)m->vcolor[4 * (size_t)id[j] + k];
                base[k] *= clamp01(v / 255.0f);
            }
        if (mp->n) {
            float wx[3], wy[3];
            pbary(s, Pcx + 256, Pcy, wx);
            pbary(s, Pcx, Pcy + 256, wy);
            float uv[2] = {0, 0}, uvx[2] = {0, 0}, uvy[2] = {0, 0};
            for (int j = 0; j < 3; ++j)
                for (int k = 0; k < 2; ++k) {
                    float q = m->uv[2 * (size_t)id[j] + k];
                    uv[k] += w[j] * q; uvx[k] += wx[j] * q; uvy[k] += wy[j] * q;
                }
            float dux = (uvx[0] - uv[0]) * tw, dvx = (uvx[1] - uv[1]) * th;
            float duy = (uvy[0] - uv[0]) * tw, dvy = (uvy[1] - uv[1]) * th;
            float rho2 = fmaxf(dux * dux + dvx * dvx, duy * duy + dvy * dvy);
            float lod = 0.5f * log2f(fmaxf(rho2, 1e-20f));
            lod = clampf(lod, 0.0f, (float)(mp->n - 1));
            int l0 = (int)floorf(lod), l1 = l0 + 1 < mp->n ? l0 + 1 : mp->n - 1;
            float fl = lod - (float)l0;
            float ta[3], tb[3];
            tex_fetch(mp, l0, uv[0], uv[1], ta);
            if (l1 != l0) {
                tex_fetch(mp, l1, uv[0], uv[1], tb);
                for (int k = 0; k < 3; ++k) ta[k] += (tb[k] - ta[k]) * fl;
            }
            for (int k = 0; k < 3; ++k) base[k] *= srgb_decode(ta[k]);
        }
    }

    float rough = clampf(roughness, 0.04f, 1.0f);
    float alpha = rough * rough, alpha2 = alpha * alpha;
    float kk = (rough + 1.0f) * (rough + 1.0f) / 8.0f;
    float G = (nl / (nl * (1.0f - kk) + kk)) * (nv / (nv * (1.0f - kk) + kk));
    float dd = nh * nh * (alpha2 - 1.0f) + 1.0f;
    float D = alpha2 / ((float)M_PI * dd * dd);
    float spec = G * D / (4.0f * nl * nv + 1e-4f);
    for (int k = 0; k < 3; ++k) {
        float f0 = 0.04f + (base[k] - 0.04f) * metallic;   /* F = f0 because v.h = 1 */
        float cdiff = base[k] * 0.96f * (1.0f - metallic);
        float col = nl * radiance * ((1.0f - f0) * cdiff / (float)M_PI + f0 * spec) + base[k] * sc->ambient;
        float enc = clamp01(powf(fmaxf(col, 0.0f), 1.0f / 2.2f));
        rgba[k] = (uint8_t)(enc * 255.0f + 0.5f);
    }
    rgba[3] = 255;
}

/* Renders one image: rgba8 (H*W*4) and depth (H*W).  zb/ib: 4*W*H scratch each. */
static void render_one(const OMesh* m, const Mips* mp, const OScene* sc, const float* T, const float* K, int W, int H,
                       uint8_t* rgba, float* depth, uint32_t* zb, uint32_t* ib) {
    const size_t np = (size_t)W * H;
    for (size_t i = 0; i < 4 * np; ++i) { zb[i] = 0xFFFFFFu; ib[i] = 0xFFFFFFFFu; }
    const double n = sc->znear, f = sc->zfar;
    const float c0 = (float)(f / (f - n)), c1 = (float)(f * n / (f - n));
    Setup s;
    for (int t = 0; t < m->T; ++t) {
        if (!tri_setup(m, t, T, K, sc->znear, &s)) continue;
        int xmin = s.X[0], xmax = s.X[0], ymin = s.Y[0], ymax = s.Y[0];
        for (int k = 1; k < 3; ++k) {
            if (s.X[k] < xmin) xmin = s.X[k];
            if (s.X[k] > xmax) xmax = s.X[k];
            if (s.Y[k] < ymin) ymin = s.Y[k];
            if (s.Y[k] > ymax) ymax = s.Y[k];
        }
        int px0 = cdiv256(xmin - 224), px1 = fdiv256(xmax - 32), py0 = cdiv256(ymin - 224), py1 = fdiv256(ymax - 32);
        if (px0 < 0) px0 = 0;
        if (py0 < 0) py0 = 0;
        if (px1 > W - 1) px1 = W - 1;
        if (py1 > H - 1) py1 = H - 1;
        if (px0 > px1 || py0 > py1) continue;
        int dx[3], dy[3], tl[3];
        for (int k = 0; k < 3; ++k) {
            int j = (k + 1) % 3;
            dx[k] = s.X[j] - s.X[k];
            dy[k] = s.Y[j] - s.Y[k];
            tl[k] = (dy[k] < 0) || (dy[k] == 0 && dx[k] > 0);   /* top-left fill rule, v down */
        }
        for (int py = py0; py <= py1; ++py)
            for (int px = px0; px <= px1; ++px)
                for (int q = 0; q < 4; ++q) {
                    int Px = px * 256 + SMP_X[q], Py = py * 256 + SMP_Y[q];
                    int inside = 1;
                    for (int k = 0; k < 3 && inside; ++k) {
                        int64_t e = (int64_t)dx[k] * (Py - s.Y[k]) - (int64_t)dy[k] * (Px - s.X[k]);
                        if (e < 0 || (e == 0 && !tl[k])) inside = 0;
                    }
                    if (!inside) continue;
                    float iz = fmaf(s.gy, (float)(Py - s.Y[0]), fmaf(s.gx, (float)(Px - s.X[0]), s.iz0));
                    float d = fmaf(-c1, iz, c0);
                    if (!(d >= 0.0f && d <= 1.0f)) continue;            /* near/far clip */
                    uint32_t d24 = (uint32_t)lrintf(d * 16777215.0f);
                    size_t si = (size_t)q * np + (size_t)py * W + px;
                    if (d24 < zb[si]) { zb[si] = d24; ib[si] = (uint32_t)t; }   /* GL_LESS, in-order */
                }
    }
    /* resolve (R5) + depth read-back (R6) */
    const float rb_num = (float)(2.0 * n * f), rb_sum = (float)(f + n), rb_diff = (float)(f - n);
    for (int py = 0; py < H; ++py)
        for (int px = 0; px < W; ++px) {
            size_t p = (size_t)py * W + px;
            unsigned sum[4] = {0, 0, 0, 0};
            uint32_t seen_t[4];
            uint8_t seen_c[4][4];
            int nseen = 0;
            for (int q = 0; q < 4; ++q) {
                uint32_t t = ib[(size_t)q * np + p];
                if (t == 0xFFFFFFFFu) continue;
                int hit = -1;
                for (int j = 0; j < nseen; ++j)
                    if (seen_t[j] == t) hit = j;
                if (hit < 0) {
                    tri_setup(m, (int)t, T, K, sc->znear, &s);
                    shade(m, mp, sc, T, K, (int)t, &s, px, py, seen_c[nseen]);
                    seen_t[nseen] = t;
                    hit = nseen++;
                }
                for (int k = 0; k < 4; ++k) sum[k] += seen_c[hit][k];
            }
            for (int k = 0; k < 4; ++k) rgba[4 * p + k] = (uint8_t)((sum[k] + 2) >> 2);
            uint32_t d24 = zb[p];   /* sample 0 */
            if (d24 == 0xFFFFFFu) depth[p] = 0.0f;
            else {
                float d = (float)d24 / 16777215.0f;
                float zn = 2.0f * d - 1.0f;
                float tmp = zn * rb_diff;
                depth[p] = rb_num / (rb_sum - tmp);
            }
        }
}

static void scene_from(const float* sp, OScene* sc) {
    sc->ambient = sp[0]; sc->intensity = sp[1]; sc->inner_cone = sp[2]; sc->outer_cone = sp[3];
    sc->znear = sp[4]; sc->zfar = sp[5];
}

/*
 * Renders B poses of ONE mesh.  Outputs (any may be NULL):
 *   rgb  (B,H,W,3) f32 max-normalised (renderer.py:103-104; data_loaders.py:112-114)
 *   depth(B,H,W)   f32
 *   rgba8(B,H,W,4) u8 resolved framebuffer before normalisation
 * material: base_color[4], metallic, roughness.  scene: ambient,intensity,inner,outer,znear,zfar.
 * nthreads <= 0: all available.
 */
static int render_batch_common(OMesh m, const float* scene, const float* T_CO, const float* K, int B, int W, int H,
                               float* rgb, float* depth, uint8_t* rgba8, int nthreads);

EXPORT int oracle_render_batch(const float* verts, int V, const int32_t* faces, int Tn, const float* uv,
                               const uint8_t* vcolor, const uint8_t* tex, int tw, int th, const float* material,
                               const float* scene, const float* T_CO, const float* K, int B, int W, int H,
                               float* rgb, float* depth, uint8_t* rgba8, int nthreads) {
    OMesh m;
    memset(&m, 0, sizeof(m));
    m.verts = verts; m.faces = faces; m.V = V; m.T = Tn; m.uv = uv; m.vcolor = vcolor; m.tex = tex; m.tw = tw; m.th = th;
    for (int k = 0; k < 4; ++k) m.base_color[k] = material[k];
    m.metallic = material[4]; m.roughness = material[5];
    return render_batch_common(m, scene, T_CO, K, B, W, H, rgb, depth, rgba8, nthreads);
}

/* Multi-part variant: face_part (T), part_material (nparts*6), part_tex[k] = RGB8 (part_th[k], part_tw[k], 3) or NULL,
 * part_use_vcolor (nparts). */
EXPORT int oracle_render_batch_parts(const float* verts, int V, const int32_t* faces, int Tn, const float* uv,
                                     const uint8_t* vcolor, const uint8_t* face_part, int nparts, const float* part_material,
                                     const uint8_t* const* part_tex, const int* part_tw, const int* part_th,
                                     const int* part_use_vcolor, const int* part_double_sided, const float* scene, const float* T_CO, const float* K,
                                     int B, int W, int H, float* rgb, float* depth, uint8_t* rgba8, int nthreads) {
    OMesh m;
    memset(&m, 0, sizeof(m));
    m.verts = verts; m.faces = faces; m.V = V; m.T = Tn; m.uv = uv; m.vcolor = vcolor;
    m.face_part = face_part; m.nparts = nparts; m.part_material = part_material; m.part_use_vcolor = part_use_vcolor; m.part_double_sided = part_double_sided;
    m.part_tw = part_tw; m.part_th = part_th;
    Mips* pm = (Mips*)calloc((size_t)nparts, sizeof(Mips));
    if (!pm) return -1;
    for (int k = 0; k < nparts; ++k) mips_build(&pm[k], (part_tex[k] && uv) ? part_tex[k] : NULL, part_tw[k], part_th[k]);
    m.part_mips = pm;
    int rc = render_batch_common(m, scene, T_CO, K, B, W, H, rgb, depth, rgba8, nthreads);
    for (int k = 0; k < nparts; ++k) mips_free(&pm[k]);
    free(pm);
    return rc;
}

static int render_batch_common(OMesh m, const float* scene, const float* T_CO, const float* K, int B, int W, int H,
                               float* rgb, float* depth, uint8_t* rgba8, int nthreads) {
    OScene sc;
    scene_from(scene, &sc);
    Mips mp;
    mips_build(&mp, (m.tex && m.uv) ? m.tex : NULL, m.tw, m.th);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    const size_t np = (size_t)W * H;
    int fail = 0;
#pragma omp parallel
    {
        uint32_t* zb = (uint32_t*)malloc(sizeof(uint32_t) * 4 * np);
        uint32_t* ib = (uint32_t*)malloc(sizeof(uint32_t) * 4 * np);
        uint8_t* fb = (uint8_t*)malloc(4 * np);
        float* db = (float*)malloc(sizeof(float) * np);
        if (!zb || !ib || !fb || !db) {
#pragma omp atomic write
            fail = 1;
        } else {
#pragma omp for schedule(dynamic, 1)
            for (int b = 0; b < B; ++b) {
                render_one(&m, &mp, &sc, T_CO + 16 * (size_t)b, K + 9 * (size_t)b, W, H, fb, db, zb, ib);
                if (rgba8) memcpy(rgba8 + 4 * np * b, fb, 4 * np);
                if (depth) memcpy(depth + np * b, db, sizeof(float) * np);
                if (rgb) {
                    /* renderer.py:82 color/255.0 (f64); :103-104 img/max; data_loaders.py:114 astype f32 */
                    unsigned qmax = 0;
                    for (size_t i = 0; i < np; ++i)
                        for (int k = 0; k < 3; ++k)
                            if (fb[4 * i + k] > qmax) qmax = fb[4 * i + k];
                    float* o = rgb + 3 * np * b;
                    for (size_t i = 0; i < np; ++i)
                        for (int k = 0; k < 3; ++k) {
                            double v = (double)fb[4 * i + k] / 255.0;
                            if (qmax > 0) v = v / ((double)qmax / 255.0);
                            o[3 * i + k] = (float)v;
                        }
                }
            }
        }
        free(zb); free(ib); free(fb); free(db);
    }
    mips_free(&mp);
    return fail ? -1 : 0;
}

/*
 * data_loaders.py:158-177 prepare_model_input (with normalize_img :79-82 and normalize_depth :69-77).
 * init (B,H,W,3) f32, gt (B,H,W,3) f32, depth (B,H,W) f32 -> x (B,C,H,W) f32, C = 6 or 7.
 * The numpy restatement in oracle/assemble_ref.py is the checker of this C version (used for timing).
 */
EXPORT int oracle_prepare_model_input(const float* init, const float* gt, const float* depth, int B, int W, int H,
                                      int use_norm_depth, float* x, int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    const size_t np = (size_t)W * H;
    const int C = use_norm_depth ? 7 : 6;
#pragma omp parallel for schedule(static)
    for (int b = 0; b < B; ++b) {
        const float* gi = gt + 3 * np * b;
        const float* ii = init + 3 * np * b;
        float* xo = x + (size_t)C * np * b;
        int ch = 0;
        if (use_norm_depth) {
            const float* d = depth + np * b;
            float dmax = 0.0f;
            double s = 0.0, ss = 0.0;
            size_t cnt = 0;
            for (size_t i = 0; i < np; ++i) {
                if (d[i] > dmax) dmax = d[i];
                if (d[i] > 0.01f) { s += d[i]; cnt++; }
            }
            int passthrough = !(dmax > 0.1f) || cnt == 0;
            float mean = 0.0f, sd = 1.0f;
            if (!passthrough) {
                double mu = s / (double)cnt;
                for (size_t i = 0; i < np; ++i)
                    if (d[i] > 0.01f) { double e = d[i] - mu; ss += e * e; }
                mean = (float)mu;
                sd = (float)sqrt(ss / (double)cnt);
                if (fabsf(sd) <= 1e-8f) passthrough = 1;   /* np.isclose(std, 0.0) */
            }
            for (size_t i = 0; i < np; ++i)
                xo[i] = passthrough ? d[i] : (d[i] > 0.01f ? (d[i] - mean) / sd : 0.0f);
            ch = 1;
        }
        for (size_t i = 0; i < np; ++i)
            for (int k = 0; k < 3; ++k) xo[(size_t)(ch + k) * np + i] = ii[3 * i + k];
        double s = 0.0, ss = 0.0;
        for (size_t i = 0; i < 3 * np; ++i) s += gi[i];
        double mu = s / (double)(3 * np);
        for (size_t i = 0; i < 3 * np; ++i) { double e = gi[i] - mu; ss += e * e; }
        double sd = sqrt(ss / (double)(3 * np));
        for (size_t i = 0; i < np; ++i)
            for (int k = 0; k < 3; ++k) xo[(size_t)(ch + 3 + k) * np + i] = (float)(((double)gi[3 * i + k] - mu) / sd);
    }
    return 0;
}

EXPORT int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
