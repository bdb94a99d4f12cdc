// rac_api.cu — library state: error string, launch counter, device mesh table.
// ref: renderer.py:28-48 (add_object) is what rac_mesh_create replaces: the
// reference re-parses and re-uploads the mesh for every rendered image; here a
// mesh is uploaded once and lives in HBM behind an integer handle.
#include "rac_common.cuh"
#include <mutex>
#include <vector>
#include <string.h>
#include <math.h>

namespace rac {

static thread_local char tl_error[512] = "";
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(tl_error, sizeof(tl_error), fmt, ap);
    va_end(ap);
}

struct MeshHost {
    bool used = false;
    int device = -1;
    int V = 0, T = 0;
    void* d_block = nullptr;  // one allocation holding every array of the mesh
};

static std::mutex g_mu;
static MeshHost g_meshes[kMaxMeshes];
static MeshDev* g_table_dev = nullptr;
static int g_table_device = -1;

const MeshDev* mesh_table_dev() { return g_table_dev; }

int mesh_host_info(rac_mesh_t h, int* V, int* T) {
    if (h < 0 || h >= kMaxMeshes || !g_meshes[h].used) return -1;
    if (V) *V = g_meshes[h].V;
    if (T) *T = g_meshes[h].T;
    return 0;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// sRGB-encoded RGBA8 box-filtered mip chain (GL generates mips on the encoded values for a
// non-sRGB internal format, which is what pyrender uploads: GL_RGBA, not GL_SRGB8_ALPHA8).
static void build_mips(const uint8_t* rgb, int tw, int th, std::vector<uchar4>& out, int& nmips) {
    std::vector<uchar4> cur((size_t)tw * th);
    for (size_t i = 0; i < (size_t)tw * th; ++i)
        cur[i] = make_uchar4(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], 255);
    nmips = 0;
    int w = tw, h = th;
    for (;;) {
        out.insert(out.end(), cur.begin(), cur.end());
        ++nmips;
        if (w == 1 && h == 1) break;
        int nw = w > 1 ? w / 2 : 1, nh = h > 1 ? h / 2 : 1;
        std::vector<uchar4> nxt((size_t)nw * nh);
        for (int y = 0; y < nh; ++y)
            for (int x = 0; x < nw; ++x) {
                int x0 = (2 * x < w) ? 2 * x : w - 1, x1 = (2 * x + 1 < w) ? 2 * x + 1 : w - 1;
                int y0 = (2 * y < h) ? 2 * y : h - 1, y1 = (2 * y + 1 < h) ? 2 * y + 1 : h - 1;
                const uchar4 a = cur[(size_t)y0 * w + x0], b = cur[(size_t)y0 * w + x1];
                const uchar4 c = cur[(size_t)y1 * w + x0], d = cur[(size_t)y1 * w + x1];
                nxt[(size_t)y * nw + x] = make_uchar4((a.x + b.x + c.x + d.x + 2) >> 2,
                                                      (a.y + b.y + c.y + d.y + 2) >> 2,
                                                      (a.z + b.z + c.z + d.z + 2) >> 2, 255);
            }
        cur.swap(nxt);
        w = nw;
        h = nh;
    }
}

}  // namespace rac

using namespace rac;

extern "C" {

const char* rac_last_error(void) { return tl_error; }
int rac_version(void) { return 100; }
long long rac_launch_count(void) { return g_launches.load(); }

void rac_scene_defaults(rac_scene_t* s) {
    if (!s) return;
    s->ambient = 0.1f;                         // ref: renderer.py:97
    s->light_intensity = 15.0f;                // ref: renderer.py:54
    s->inner_cone = (float)(M_PI / 9.0);       // ref: renderer.py:55
    s->outer_cone = (float)(M_PI / 2.0);       // ref: renderer.py:56
    s->znear = 0.05f;                          // pyrender DEFAULT_Z_NEAR
    s->zfar = 100.0f;                          // pyrender DEFAULT_Z_FAR
}

void rac_material_defaults(rac_material_t* m) {
    if (!m) return;
    m->base_color[0] = m->base_color[1] = m->base_color[2] = 0.3f;
    m->base_color[3] = 1.0f;
    m->metallic = 0.2f;
    m->roughness = 0.8f;
}

int rac_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    RAC_CUDA_CHECK(cudaGetDevice(&dev));
    cudaDeviceProp p;
    RAC_CUDA_CHECK(cudaGetDeviceProperties(&p, dev));
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    return RAC_OK;
}

// Shared body of rac_mesh_create / rac_mesh_create_parts.  n_parts == 0: single material `mat`, texture h_tex.
static int mesh_create_common(const float* h_verts, int V, const int32_t* h_faces, int T,
                              const float* h_uv, const uint8_t* h_tex, int tw, int th,
                              const uint8_t* h_vcolor, const rac_material_t* mat,
                              const uint8_t* h_face_part, int n_parts, const rac_part_t* parts, rac_mesh_t* out) {
    RAC_REQUIRE(h_verts && h_faces && out, "rac_mesh_create: null verts/faces/out");
    RAC_REQUIRE(V > 0 && T > 0, "rac_mesh_create: V=%d T=%d must be positive", V, T);
    RAC_REQUIRE(!h_tex || (h_uv && tw > 0 && th > 0), "rac_mesh_create: texture needs uv and tw,th>0");
    for (int i = 0; i < 3 * T; ++i)
        RAC_REQUIRE(h_faces[i] >= 0 && h_faces[i] < V, "rac_mesh_create: face index %d out of range", h_faces[i]);
    if (n_parts > 0) {
        RAC_REQUIRE(h_face_part && parts && n_parts <= 255, "rac_mesh_create_parts: needs face_part, parts, n_parts <= 255");
        for (int t = 0; t < T; ++t)
            RAC_REQUIRE(h_face_part[t] < n_parts, "rac_mesh_create_parts: face %d names part %d of %d", t, (int)h_face_part[t], n_parts);
        for (int k = 0; k < n_parts; ++k)
            RAC_REQUIRE(!parts[k].tex_rgb8 || (h_uv && parts[k].tex_w > 0 && parts[k].tex_h > 0),
                        "rac_mesh_create_parts: part %d has a texture but no uv / size", k);
    }

    int dev = 0;
    RAC_CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_table_dev) {
        RAC_CUDA_CHECK(cudaMalloc(&g_table_dev, sizeof(MeshDev) * kMaxMeshes));
        RAC_CUDA_CHECK(cudaMemset(g_table_dev, 0, sizeof(MeshDev) * kMaxMeshes));
        g_table_device = dev;
    }
    RAC_REQUIRE(dev == g_table_device, "rac_mesh_create: one process drives one GPU (table on %d, current %d)",
                g_table_device, dev);
    int slot = -1;
    for (int i = 0; i < kMaxMeshes; ++i)
        if (!g_meshes[i].used) { slot = i; break; }
    RAC_REQUIRE(slot >= 0, "rac_mesh_create: mesh table full (%d)", kMaxMeshes);

    // face normals: normalize((v1-v0)x(v2-v0)), zero for degenerate faces (trimesh convention, R3)
    std::vector<float> fn((size_t)3 * T);
    for (int t = 0; t < T; ++t) {
        const float* a = h_verts + 3 * (size_t)h_faces[3 * t];
        const float* b = h_verts + 3 * (size_t)h_faces[3 * t + 1];
        const float* c = h_verts + 3 * (size_t)h_faces[3 * t + 2];
        double e1[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]};
        double e2[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
        double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        double len = sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        for (int k = 0; k < 3; ++k) fn[3 * (size_t)t + k] = len > 0.0 ? (float)(n[k] / len) : 0.0f;
    }
    // mip chains: the mesh's own texture first, then one chain per textured part
    std::vector<uchar4> mips;
    int nmips = 0;
    if (h_tex) build_mips(h_tex, tw, th, mips, nmips);
    std::vector<size_t> part_mip_off((size_t)(n_parts > 0 ? n_parts : 0), 0);
    std::vector<int> part_nmips((size_t)(n_parts > 0 ? n_parts : 0), 0);
    for (int k = 0; k < n_parts; ++k)
        if (parts[k].tex_rgb8) {
            part_mip_off[k] = mips.size();
            std::vector<uchar4> chain;
            build_mips(parts[k].tex_rgb8, parts[k].tex_w, parts[k].tex_h, chain, part_nmips[k]);
            mips.insert(mips.end(), chain.begin(), chain.end());
        }

    const size_t o_verts = 0;
    const size_t o_faces = align_up(o_verts + sizeof(float) * 3 * V, 256);
    const size_t o_fn = align_up(o_faces + sizeof(int) * 3 * T, 256);
    const size_t o_uv = align_up(o_fn + sizeof(float) * 3 * T, 256);
    const size_t o_vc = align_up(o_uv + (h_uv ? sizeof(float) * 2 * V : 0), 256);
    const size_t o_tex = align_up(o_vc + (h_vcolor ? sizeof(uchar4) * V : 0), 256);
    const size_t o_parts = align_up(o_tex + sizeof(uchar4) * mips.size(), 256);
    const size_t o_fpart = align_up(o_parts + sizeof(PartDev) * (size_t)(n_parts > 0 ? n_parts : 0), 256);
    const size_t total = align_up(o_fpart + (n_parts > 0 ? (size_t)T : 0), 256);

    char* blk = nullptr;
    RAC_CUDA_CHECK(cudaMalloc(&blk, total));
    std::vector<PartDev> pd((size_t)(n_parts > 0 ? n_parts : 0));
    for (int k = 0; k < n_parts; ++k) {
        PartDev& q = pd[k];
        memset(&q, 0, sizeof(q));
        for (int c = 0; c < 4; ++c) q.base_color[c] = parts[k].material.base_color[c];
        q.metallic = parts[k].material.metallic;
        const float rough = fminf(fmaxf(parts[k].material.roughness, 0.04f), 1.0f);     // (same formulas as make_shade_ctx)
        const float alpha = rough * rough;
        q.alpha2 = alpha * alpha;
        q.kgeo = (rough + 1.0f) * (rough + 1.0f) * 0.125f;
        q.has_vcolor = parts[k].use_vertex_color ? 1 : 0;
        q.double_sided = parts[k].double_sided ? 1 : 0;
        if (parts[k].tex_rgb8) {
            q.tex = (const uchar4*)(blk + o_tex) + part_mip_off[k];
            q.tw = parts[k].tex_w; q.th = parts[k].tex_h; q.nmips = part_nmips[k];
        }
    }
    cudaError_t e = cudaMemcpy(blk + o_verts, h_verts, sizeof(float) * 3 * V, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(blk + o_faces, h_faces, sizeof(int) * 3 * T, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(blk + o_fn, fn.data(), sizeof(float) * 3 * T, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && h_uv) e = cudaMemcpy(blk + o_uv, h_uv, sizeof(float) * 2 * V, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && h_vcolor) e = cudaMemcpy(blk + o_vc, h_vcolor, sizeof(uchar4) * V, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !mips.empty())
        e = cudaMemcpy(blk + o_tex, mips.data(), sizeof(uchar4) * mips.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n_parts > 0) e = cudaMemcpy(blk + o_parts, pd.data(), sizeof(PartDev) * n_parts, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n_parts > 0) e = cudaMemcpy(blk + o_fpart, h_face_part, (size_t)T, cudaMemcpyHostToDevice);

    MeshDev md;
    memset(&md, 0, sizeof(md));
    md.verts = (const float*)(blk + o_verts);
    md.faces = (const int*)(blk + o_faces);
    md.fnormals = (const float*)(blk + o_fn);
    md.uv = h_uv ? (const float*)(blk + o_uv) : nullptr;
    md.vcolor = h_vcolor ? (const uchar4*)(blk + o_vc) : nullptr;
    md.tex = h_tex ? (const uchar4*)(blk + o_tex) : nullptr;
    md.V = V;
    md.T = T;
    md.tw = h_tex ? tw : 0;
    md.th = h_tex ? th : 0;
    md.nmips = nmips;
    md.parts = n_parts > 0 ? (const PartDev*)(blk + o_parts) : nullptr;
    md.face_part = n_parts > 0 ? (const unsigned char*)(blk + o_fpart) : nullptr;
    rac_material_t m;
    rac_material_defaults(&m);
    if (mat) m = *mat;
    for (int k = 0; k < 4; ++k) md.base_color[k] = m.base_color[k];
    md.metallic = m.metallic;
    md.roughness = m.roughness;
    if (e == cudaSuccess) e = cudaMemcpy(g_table_dev + slot, &md, sizeof(md), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(blk);
        set_error("rac_mesh_create: upload failed: %s", cudaGetErrorString(e));
        return RAC_E_CUDA;
    }
    g_meshes[slot].used = true;
    g_meshes[slot].device = dev;
    g_meshes[slot].V = V;
    g_meshes[slot].T = T;
    g_meshes[slot].d_block = blk;
    *out = slot;
    return RAC_OK;
}

int rac_mesh_create(const float* h_verts, int V, const int32_t* h_faces, int T,
                    const float* h_uv, const uint8_t* h_tex, int tw, int th,
                    const uint8_t* h_vcolor, const rac_material_t* mat, rac_mesh_t* out) {
    return mesh_create_common(h_verts, V, h_faces, T, h_uv, h_tex, tw, th, h_vcolor, mat, nullptr, 0, nullptr, out);
}

int rac_mesh_create_parts(const float* h_verts, int V, const int32_t* h_faces, int T,
                          const float* h_uv, const uint8_t* h_vcolor,
                          const uint8_t* h_face_part, int n_parts, const rac_part_t* parts, rac_mesh_t* out) {
    RAC_REQUIRE(n_parts > 0, "rac_mesh_create_parts: n_parts=%d must be positive", n_parts);
    return mesh_create_common(h_verts, V, h_faces, T, h_uv, nullptr, 0, 0, h_vcolor, nullptr, h_face_part, n_parts, parts, out);
}

int rac_mesh_destroy(rac_mesh_t mesh) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (mesh < 0 || mesh >= kMaxMeshes || !g_meshes[mesh].used) {
        set_error("rac_mesh_destroy: unknown handle %d", mesh);
        return RAC_E_HANDLE;
    }
    // the mesh may still be read by kernels in flight on any stream
    RAC_CUDA_CHECK(cudaDeviceSynchronize());
    RAC_CUDA_CHECK(cudaFree(g_meshes[mesh].d_block));
    MeshDev zero;
    memset(&zero, 0, sizeof(zero));
    RAC_CUDA_CHECK(cudaMemcpy(g_table_dev + mesh, &zero, sizeof(zero), cudaMemcpyHostToDevice));
    g_meshes[mesh] = MeshHost();
    return RAC_OK;
}

int rac_mesh_info(rac_mesh_t mesh, int* V, int* T) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (mesh_host_info(mesh, V, T) != 0) {
        set_error("rac_mesh_info: unknown handle %d", mesh);
        return RAC_E_HANDLE;
    }
    return RAC_OK;
}

}  // extern "C"
