// raster.cu — step 1 (+2) of the hot path: batched mesh rasterisation for sm_100a.
//
// Replaces: data_loaders.py:84-118 (render_batch) -> renderer.py:85-105 (render_scene) ->
// pyrender/OpenGL (R1-R9 in SURVEY.md §8a), and data_loaders.py:158-177 (prepare_model_input).
//
// Kernels of one call (see raster.cuh for the spec they implement):
//   setup_kernel     (raster_setup.inl) 128 threads x 2 triangles of one image: transform, project, snap, cull, 1/z
//                    plane, camera-frame normal -> 48-B record.  The CTA bins its survivors by 32x16-px bin in shared
//                    memory and publishes, without any global atomic, a private ref region (u8 local ids grouped by
//                    bin) plus one descriptor per bin; triangles spanning more than 2x2 bins — or whose UNCLIPPED extent
//                    exceeds 64 px — go to a per-image large list (int64 edge functions).  The CTA that finishes an
//                    image last appends the image's non-empty bins to 16 load-class lists.
//   raster_kernel    (raster_tile.inl, raster_shade.inl) persistent CTAs (SMs x resident CTAs) pop bins heaviest class
//                    first; every ticket comes from one atomic counter.  Per bin the 4x-MSAA visibility buffer lives in
//                    shared memory as 32-bit keys (depth24 << 8 | batch-local triangle id, 8 KB): triangles are
//                    gathered through the descriptors and scattered into it with ONE native shared-memory atomic-min per
//                    covered sample (exact int32 edge functions for binned triangles, int64 for large ones); more than
//                    254 triangles in a bin commit (shade + park) between batches.  Then every pixel is shaded + resolved
//                    once and written as RGBA8 + f32 depth; per-image max / depth statistics are reduced per CTA and
//                    flushed with atomics.
//   gt_norm_kernel   (raster_finalize.inl) on a forked stream, resident next to the raster CTAs: zero background of the
//                    render planes + normalised gt planes of x (8-CTA cluster per image, DSMEM reduction).
//   finalize_live_kernel  non-empty bins only: RGBA8 + depth (L2-resident) -> normalised f32 outputs with 128-bit
//                    coalesced stores (NCHW concat or HWC render).
//   gt_norm_tma_kernel / assemble_render_kernel / depth_stats_kernel: rac_assemble_input (existing renders).
#include "raster.cuh"
#include <cooperative_groups.h>
#include <mutex>
#include <type_traits>
#include <math.h>
#include <stdlib.h>

namespace rac {

// Diagnostic builds (-DRAC_PHASE_TIMING): per-kernel [first CTA start, last CTA end] in globaltimer ns, read back with
// rac_debug_timeline().  Kernel ids: 0 gt_norm, 1 setup, 2 raster, 3 finalize_live.
#ifdef RAC_PHASE_TIMING
__device__ unsigned long long g_tl[16];
#define RAC_TL(k)                                                                  \
    do {                                                                           \
        if (threadIdx.x == 0) {                                                    \
            unsigned long long _t;                                                 \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_t));                 \
            atomicMin(&g_tl[2 * (k)], _t);                                         \
            atomicMax(&g_tl[2 * (k) + 1], _t);                                     \
        }                                                                          \
    } while (0)
#else
#define RAC_TL(k) do { } while (0)
#endif

#include "raster_setup.inl"
#include "raster_shade.inl"
#include "raster_finalize.inl"
#include "raster_tile.inl"

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static TileGrid make_tile_grid(int W, int H, int maxT) {
    TileGrid tg;
    tg.ntx = (W + kTileW - 1) / kTileW;
    tg.nty = (H + kTileH - 1) / kTileH;
    tg.nt = tg.ntx * tg.nty;
    tg.binned = tg.nt <= kMaxBinTiles ? 1 : 0;
    tg.ncta_max = (maxT + kSetupThreads - 1) / kSetupThreads;
    return tg;
}

constexpr int kMaxRasterGrid = 256 * 4;     // upper bound of the persistent grid (SMs x resident CTAs)

static RenderWs carve_ws(void* base, int B, int W, int H, int maxT) {
    RenderWs w;
    char* p = (char*)base;
    size_t off = 0;
    const TileGrid tg = make_tile_grid(W, H, maxT);
    const size_t ndesc = tg.binned ? (size_t)B * tg.nt * tg.ncta_max : 0;
    const size_t nrefs = tg.binned ? ((size_t)B * tg.ncta_max) << 10 : 0;
    // [stats | ctl | loads] form one contiguous block that a single memset zeroes per call
    w.stats = (ImgStats*)(p + off);          off = align256(off + sizeof(ImgStats) * (size_t)B);
    w.ctl = (WorkCtl*)(p + off);             off = align256(off + sizeof(WorkCtl));
    w.loads = (unsigned*)(p + off);          off = align256(off + sizeof(unsigned) * (size_t)B * tg.nt);
    w.zero_bytes = off;
    w.buckets = (unsigned*)(p + off);        off = align256(off + sizeof(unsigned) * (size_t)kBuckets * B * tg.nt);
    w.mesh_ids = (int*)(p + off);            off = align256(off + sizeof(int) * (size_t)B);
    w.large_cnt = (unsigned*)(p + off);      off = align256(off + sizeof(unsigned) * (size_t)B * tg.ncta_max);
    w.large_tri = (unsigned*)(p + off);      off = align256(off + sizeof(unsigned) * (size_t)B * maxT);
    w.recs = (int4*)(p + off);               off = align256(off + sizeof(int4) * 3 * (size_t)B * maxT);
    w.desc = (unsigned*)(p + off);           off = align256(off + sizeof(unsigned) * ndesc);
    w.refs = (unsigned char*)(p + off);      off = align256(off + nrefs);
    w.rgba = (unsigned*)(p + off);           off = align256(off + sizeof(unsigned) * (size_t)B * W * H);
    w.depth = (float*)(p + off);             off = align256(off + sizeof(float) * (size_t)B * W * H);
    w.total = off;
    return w;
}

static SceneConsts make_scene_consts(const rac_scene_t* scene) {
    rac_scene_t s;
    rac_scene_defaults(&s);
    if (scene) s = *scene;
    SceneConsts c;
    c.znear = s.znear; c.zfar = s.zfar;
    const double n = (double)s.znear, f = (double)s.zfar;
    c.c0 = (float)(f / (f - n));
    c.c1 = (float)(f * n / (f - n));
    c.rb_num = (float)(2.0 * n * f);
    c.rb_sum = (float)(f + n);
    c.rb_diff = (float)(f - n);
    // pyrender SpotLight uniforms: light_angle_scale / light_angle_offset (R7)
    const double las = 1.0 / fmax(0.001, cos((double)s.inner_cone) - cos((double)s.outer_cone));
    c.las = (float)las;
    c.lao = (float)(-cos((double)s.outer_cone) * las);
    c.intensity = s.light_intensity;
    c.ambient = s.ambient;
    return c;
}

// Per-image mesh-id table of a mixed-mesh batch: handed to the device BY VALUE, as kernel parameters (512 ids per
// launch).  Nothing host-side outlives the call — no staging buffer to pool, reuse or leak, and a captured call keeps
// its ids inside the graph's kernel node (a memcpy node would keep a pointer into caller memory instead).
struct IdChunk { int v[512]; };
__global__ void mesh_ids_kernel(int* __restrict__ dst, const __grid_constant__ IdChunk c, int n) {
    const int i = (int)threadIdx.x;
    if (i < n) dst[i] = c.v[i];
}
static std::mutex g_stage_mu;

static int upload_ids(const int32_t* h_ids, int B, int* d_dst, cudaStream_t stream) {
    for (int base = 0; base < B; base += 512) {
        IdChunk c;
        const int n = B - base < 512 ? B - base : 512;
        for (int i = 0; i < n; ++i) c.v[i] = h_ids[base + i];
        for (int i = n; i < 512; ++i) c.v[i] = 0;
        mesh_ids_kernel<<<1, 512, 0, stream>>>(d_dst + base, c, n);
        RAC_LAUNCHED();
    }
    return RAC_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Fork / join plumbing for the gt-normalise kernel: one library-owned non-blocking stream and two events per device.
// Works the same in eager mode and under stream capture (the side stream joins the capture through the fork event
// and is joined back before render_common returns).
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    int n_sm = 0;              // per device: a process may drive several GPUs
    int raster_grid = 0;       // persistent raster CTAs (SMs x resident CTAs)
};
static SideStream g_side[64];

// Per-device state of the render path (created on first use, under g_stage_mu): the side stream and its events, the
// SM count, the persistent raster grid, and the L1 / shared-memory carve-out of every kernel of the call (an SM does
// not mix carve-outs, so the forked gt kernel and the main chain ask for raster_kernel's split).
static int side_stream(SideStream** out) {
    int dev = 0;
    RAC_CUDA_CHECK(cudaGetDevice(&dev));
    RAC_REQUIRE(dev >= 0 && dev < 64, "render: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_stage_mu);
    SideStream& s = g_side[dev];
    if (!s.stream) {
        int per_sm = 0;
        RAC_CUDA_CHECK(cudaDeviceGetAttribute(&s.n_sm, cudaDevAttrMultiProcessorCount, dev));
        RAC_CUDA_CHECK(cudaFuncSetAttribute(raster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRasterSmemBytes));
        RAC_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, raster_kernel, kRasterThreads, kRasterSmemBytes));
        if (per_sm < 1) per_sm = 1;
        s.raster_grid = s.n_sm * per_sm;
        RAC_CUDA_CHECK(cudaFuncSetAttribute(setup_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        RAC_CUDA_CHECK(cudaFuncSetAttribute(finalize_live_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        RAC_CUDA_CHECK(cudaFuncSetAttribute(gt_norm_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        RAC_CUDA_CHECK(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
        RAC_CUDA_CHECK(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
        RAC_CUDA_CHECK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));     // (last: marks the slot ready)
    }
    *out = &s;
    return RAC_OK;
}

// `fin` arrives with the caller-facing fields set (mode, outputs, gt, use_norm_depth); the workspace-side fields are
// filled here and the whole finalize pass runs as work items of raster_kernel.
static int render_common(const int32_t* h_mesh_ids, const float* d_T_CO, const float* d_K, int B, int W, int H,
                         const rac_scene_t* scene, FinArgs fin,
                         void* d_ws, size_t ws_bytes, cudaStream_t stream) {
    const float* d_gt_hwc = fin.gt_hwc;
    RAC_REQUIRE(h_mesh_ids && d_T_CO && d_K, "render: null mesh ids / T_CO / K");
    RAC_REQUIRE(B > 0 && W > 0 && H > 0, "render: B=%d W=%d H=%d must be positive", B, W, H);
    RAC_REQUIRE(W <= 8192 && H <= 8192, "render: image %dx%d exceeds the 8192 px limit", W, H);
    RAC_REQUIRE(B <= 65535, "render: B=%d exceeds 65535 (grid.y)", B);
    RAC_REQUIRE((double)B * ((W + kTileW - 1) / kTileW) * ((H + kTileH - 1) / kTileH) < 4.0e9, "render: B x bins exceeds 2^32");
    int maxT = 0;
    bool uniform = true;
    for (int i = 0; i < B; ++i) {
        int T = 0;
        if (mesh_host_info(h_mesh_ids[i], nullptr, &T) != 0) {
            set_error("render: unknown mesh handle %d at batch index %d", h_mesh_ids[i], i);
            return RAC_E_HANDLE;
        }
        if (T > maxT) maxT = T;
        if (h_mesh_ids[i] != h_mesh_ids[0]) uniform = false;
    }
    const size_t need = rac_render_workspace_bytes(B, W, H, maxT);
    if (!d_ws || ws_bytes < need) {
        set_error("render: workspace %zu B < required %zu B", ws_bytes, need);
        return RAC_E_WORKSPACE;
    }
    RAC_REQUIRE((reinterpret_cast<uintptr_t>(d_ws) & 255u) == 0, "render: workspace must be 256-B aligned");
    RenderWs ws = carve_ws(d_ws, B, W, H, maxT);
    const SceneConsts S = make_scene_consts(scene);
    const MeshDev* table = mesh_table_dev();
    const size_t npix = (size_t)W * H;
#ifdef RAC_PHASE_TIMING   // diagnostic builds only: RAC_DEBUG_SKIP bit mask (1 side work, 2 setup, 4 raster, 8 finalize, 16 no fork, 64 gt cluster per image)
    const char* skip_env = getenv("RAC_DEBUG_SKIP");
    const int skip = skip_env ? atoi(skip_env) : 0;
#else
    constexpr int skip = 0;
#endif
    // Side work — everything that does not depend on the render: the gt half of the concat (+ zero background of the
    // render planes of x) and the zero background of the other outputs.  It runs on a forked stream under
    // setup_kernel / raster_kernel and is joined before finalize_live_kernel overwrites the non-empty tiles.
    SideStream* side = nullptr;
    {
        int rc = side_stream(&side);
        if (rc != RAC_OK) return rc;
    }
    const int n_sm = side->n_sm;
    cudaStream_t sstream = (skip & 16) ? stream : side->stream;
    const TileGrid tg = make_tile_grid(W, H, maxT);
    // Concat-only calls (no depth / HWC / RGBA8 side outputs): the gt kernel leaves the quads of non-empty bins alone,
    // so it and finalize_live_kernel write disjoint memory and the side stream is joined AFTER finalize — the gt
    // kernel, which runs about as long as setup + raster, no longer sits in front of finalize on the critical path.
    const bool late_join = d_gt_hwc && fin.x_nchw && !fin.depth_out && !fin.rgb_hwc_out && !fin.rgba_copy && !skip &&
                           W % 4 == 0 && npix % 4 == 0 && aligned16(d_gt_hwc) && aligned16(fin.x_nchw);
    auto side_work = [&]() -> int {
        if (skip & 1) return RAC_OK;
        if (d_gt_hwc) {
            // one CTA per SM (n_sm / 8 clusters), launched before setup so that it is resident next to the
            // setup / raster CTAs for the whole call and streams in the background
            int rc = launch_gt_norm(d_gt_hwc, fin.x_nchw, B, (int)npix, fin.use_norm_depth, 1, (skip & 64) ? 0 : n_sm / kGtCluster, sstream,
                                    late_join ? ws.loads : nullptr, W, tg.ntx, tg.nt);
            if (rc != RAC_OK) return rc;
        }
        if (fin.depth_out) RAC_CUDA_CHECK(cudaMemsetAsync(fin.depth_out, 0, sizeof(float) * npix * B, sstream));
        if (fin.rgb_hwc_out) RAC_CUDA_CHECK(cudaMemsetAsync(fin.rgb_hwc_out, 0, sizeof(float) * 3 * npix * B, sstream));
        if (fin.rgba_copy) RAC_CUDA_CHECK(cudaMemsetAsync(fin.rgba_copy, 0, sizeof(unsigned) * npix * B, sstream));
        return RAC_OK;
    };
    const bool fork_early = (skip & 32) != 0;
    if (fork_early && !(skip & 16)) {
        RAC_CUDA_CHECK(cudaEventRecord(side->fork, stream));
        RAC_CUDA_CHECK(cudaStreamWaitEvent(side->stream, side->fork, 0));
        int rc = side_work(); if (rc != RAC_OK) return rc;
    }
    RAC_CUDA_CHECK(cudaMemsetAsync(ws.stats, 0, ws.zero_bytes, stream));
    const int* d_ids = nullptr;
    if (!uniform) {
        int rc = upload_ids(h_mesh_ids, B, ws.mesh_ids, stream);
        if (rc != RAC_OK) return rc;
        d_ids = ws.mesh_ids;
    }
    if (!(skip & 2)) {
        dim3 grid(tg.ncta_max, B);
        setup_kernel<<<grid, kSetupCtaThreads, 0, stream>>>(table, d_ids, h_mesh_ids[0], d_T_CO, d_K, W, H, maxT, tg, S.znear,
                                                         ws.stats, ws.large_cnt, ws.large_tri, ws.recs, ws.desc, ws.refs,
                                                         ws.ctl, ws.loads, ws.buckets);
        RAC_LAUNCHED();
    }
    // fork AFTER setup: setup_kernel is bound by its own memory traffic (records, descriptors) and triples in
    // duration next to a streaming kernel, while raster_kernel is issue-bound and does not notice one
    // (profiles/kernel_timeline.py)
    if (!fork_early && !(skip & 16)) {
        RAC_CUDA_CHECK(cudaEventRecord(side->fork, stream));
        RAC_CUDA_CHECK(cudaStreamWaitEvent(side->stream, side->fork, 0));
    }
    if (!fork_early || (skip & 16)) { int rc = side_work(); if (rc != RAC_OK) return rc; }
    if (!(skip & 16)) RAC_CUDA_CHECK(cudaEventRecord(side->join, side->stream));
    if (!(skip & 4)) {
        const int grid_max = side->raster_grid;
        RAC_REQUIRE(grid_max > 0 && grid_max <= kMaxRasterGrid, "render: raster grid query failed (%d)", grid_max);
        const int grid = (int)min((long long)grid_max, (long long)B * tg.nt);
        raster_kernel<<<grid, kRasterThreads, kRasterSmemBytes, stream>>>(table, d_ids, h_mesh_ids[0], d_T_CO, d_K, W, H, maxT, tg, S,
                                                                          ws.stats, ws.large_cnt, ws.large_tri, ws.recs, ws.desc, ws.refs,
                                                                          ws.ctl, ws.buckets, (size_t)B * tg.nt, ws.rgba,
                                                                          (fin.depth_out || (fin.x_nchw && fin.use_norm_depth)) ? ws.depth : (float*)nullptr,
                                                                          (fin.x_nchw && fin.use_norm_depth) ? 1 : 0);
        RAC_LAUNCHED();
    }
    if (!(skip & 16) && !late_join) RAC_CUDA_CHECK(cudaStreamWaitEvent(stream, side->join, 0));
    if (!(skip & 8)) {
        fin.stats = ws.stats; fin.rgba = ws.rgba; fin.depth = ws.depth; fin.init_hwc = nullptr;
        fin.npix = (int)npix; fin.W = W;
        fin.vec = (W % 4 == 0 && aligned16(ws.rgba) && aligned16(ws.depth) && aligned16(fin.x_nchw) &&
                   aligned16(fin.rgb_hwc_out) && aligned16(fin.rgba_copy) && aligned16(fin.depth_out)) ? 1 : 0;
        const int grid = (int)min((long long)n_sm * 16, (long long)B * tg.nt);
        finalize_live_kernel<<<grid, kFinThreads, 0, stream>>>(fin, tg, H, ws.buckets, (size_t)B * tg.nt, ws.ctl);
        RAC_LAUNCHED();
    }
    if (late_join) RAC_CUDA_CHECK(cudaStreamWaitEvent(stream, side->join, 0));
    return RAC_OK;
}

}  // namespace rac

using namespace rac;

#ifdef RAC_PHASE_TIMING
__global__ void rac_debug_epoch_kernel() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    rac::g_tl[15] = t;
}
#endif

// ---------------------------------------------------------------------------------------------
// rac_assemble_input_static: scratch that nobody has to allocate
// ---------------------------------------------------------------------------------------------
// The per-image depth statistics that rac_assemble_input keeps in caller-provided scratch (one 64-byte ImgStats per
// image) live here in a fixed array that is part of the library's device image.  Each GPU that loads librac_b200 owns
// one copy (RAC_ASSEMBLE_STATIC_MAX_B * 64 B = 64 KiB) from module load until the process exits.
//
// In HOST code the identifier `g_assemble_static` is only a handle for the runtime's symbol table; it is not a device
// address and must never be handed to a kernel.  The address that is valid on the GPU comes from
// cudaGetSymbolAddress, asked once per device while that device is current, and is remembered per device ordinal.
static __device__ __align__(256) unsigned char g_assemble_static[RAC_ASSEMBLE_STATIC_MAX_B * sizeof(rac::ImgStats)];

constexpr int kStaticMaxDevices = 64;
static std::atomic<void*> g_assemble_static_addr[kStaticMaxDevices];

// Device address of g_assemble_static on the CURRENT device.  The first call on a device performs the symbol lookup
// (which may load the module and synchronise — hence "not inside a stream capture"); later calls are one atomic load.
static int assemble_static_scratch(void** out) {
    int dev = 0;
    RAC_CUDA_CHECK(cudaGetDevice(&dev));
    RAC_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "rac_assemble_input_static: device ordinal %d out of range", dev);
    void* p = g_assemble_static_addr[dev].load(std::memory_order_acquire);
    if (!p) {
        RAC_CUDA_CHECK(cudaGetSymbolAddress(&p, g_assemble_static));
        RAC_REQUIRE(p != nullptr, "rac_assemble_input_static: cudaGetSymbolAddress returned a null address");
        g_assemble_static_addr[dev].store(p, std::memory_order_release);   // (two racing threads store the same value)
    }
    *out = p;
    return RAC_OK;
}

extern "C" {

#ifdef RAC_PHASE_TIMING
__attribute__((visibility("default"))) int rac_debug_timeline(unsigned long long* h_out16, int reset) {
    if (h_out16) cudaMemcpyFromSymbol(h_out16, g_tl, sizeof(unsigned long long) * 16);
    if (reset) {
        unsigned long long init[16];
        for (int i = 0; i < 16; ++i) init[i] = (i & 1 || i >= 8) ? 0ull : ~0ull;
        cudaMemcpyToSymbol(g_tl, init, sizeof(init));
        rac_debug_epoch_kernel<<<1, 1>>>();
        cudaDeviceSynchronize();
    }
    return 0;
}
#endif

size_t rac_render_workspace_bytes(int B, int W, int H, int maxT) {
    if (B <= 0 || W <= 0 || H <= 0 || maxT <= 0) return 0;
    return carve_ws(nullptr, B, W, H, maxT).total;
}

size_t rac_assemble_workspace_bytes(int B) {
    return B > 0 ? align256(sizeof(ImgStats) * (size_t)B) : 0;
}

int rac_render_batch(const int32_t* h_mesh_ids, const float* d_T_CO, const float* d_K, int B, int W, int H,
                     const rac_scene_t* scene, float* d_rgb_hwc, float* d_depth, uint8_t* d_rgba8,
                     void* d_ws, size_t ws_bytes, rac_stream_t stream_) {
    FinArgs fin = {};
    fin.mode = 1;
    fin.rgb_hwc_out = d_rgb_hwc;
    fin.rgba_copy = (unsigned*)d_rgba8;
    fin.depth_out = d_depth;
    return render_common(h_mesh_ids, d_T_CO, d_K, B, W, H, scene, fin, d_ws, ws_bytes, (cudaStream_t)stream_);
}

int rac_render_concat(const int32_t* h_mesh_ids, const float* d_T_CO, const float* d_K, int B, int W, int H,
                      const rac_scene_t* scene, const float* d_gt_hwc, int use_norm_depth, float* d_x_nchw,
                      float* d_depth_out, float* d_rgb_hwc_out, void* d_ws, size_t ws_bytes, rac_stream_t stream_) {
    RAC_REQUIRE(d_gt_hwc && d_x_nchw, "rac_render_concat: null gt / x");
    FinArgs fin = {};
    fin.mode = 0;
    fin.gt_hwc = d_gt_hwc;
    fin.use_norm_depth = use_norm_depth;
    fin.x_nchw = d_x_nchw;
    fin.rgb_hwc_out = d_rgb_hwc_out;
    fin.depth_out = d_depth_out;
    return render_common(h_mesh_ids, d_T_CO, d_K, B, W, H, scene, fin, d_ws, ws_bytes, (cudaStream_t)stream_);
}

int rac_assemble_input(const float* d_init_hwc, const float* d_gt_hwc, const float* d_depth, int B, int W, int H,
                       int use_norm_depth, float* d_x_nchw, void* d_stats_ws, size_t ws_bytes, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_init_hwc && d_x_nchw, "rac_assemble_input: null init / x");
    RAC_REQUIRE(B > 0 && W > 0 && H > 0 && B <= 65535, "rac_assemble_input: bad shape B=%d W=%d H=%d", B, W, H);
    RAC_REQUIRE(!use_norm_depth || d_depth, "rac_assemble_input: use_norm_depth needs depth");
    if (!d_stats_ws || ws_bytes < rac_assemble_workspace_bytes(B)) {
        set_error("rac_assemble_input: workspace %zu B < required %zu B", ws_bytes, rac_assemble_workspace_bytes(B));
        return RAC_E_WORKSPACE;
    }
    ImgStats* stats = (ImgStats*)d_stats_ws;
    const int npix = W * H;
    RAC_CUDA_CHECK(cudaMemsetAsync(stats, 0, sizeof(ImgStats) * (size_t)B, stream));
    if (use_norm_depth) {
        const int chunks = (npix + 8191) / 8192;
        depth_stats_kernel<<<dim3(chunks, B), 256, 0, stream>>>(stats, d_depth, npix, chunks);
        RAC_LAUNCHED();
    }
    if (d_gt_hwc) {      // NULL: the caller fills the gt planes (cached across the refinement iterations of a batch)
        int rc = launch_gt_norm(d_gt_hwc, d_x_nchw, B, npix, use_norm_depth, 0, 0, stream);
        if (rc != RAC_OK) return rc;
    }
    FinArgs fin = {};
    fin.mode = 0;
    fin.stats = stats; fin.init_hwc = d_init_hwc; fin.depth = d_depth; fin.gt_hwc = d_gt_hwc;
    fin.npix = npix; fin.W = W; fin.use_norm_depth = use_norm_depth; fin.x_nchw = d_x_nchw;
    fin.vec = (npix % 4 == 0 && aligned16(d_init_hwc) && aligned16(d_depth) && aligned16(d_x_nchw)) ? 1 : 0;
    assemble_render_kernel<<<dim3((npix + kFinChunkPix - 1) / kFinChunkPix, B), 256, 0, stream>>>(fin);
    RAC_LAUNCHED();
    return RAC_OK;
}

// rac_assemble_input for callers that cannot allocate device memory (see include/rac.h).  Same kernels, same
// arithmetic; the only difference is where the ImgStats scratch lives.  Nothing in here allocates: no cudaMalloc /
// cudaFree, no stream or event creation, no torch.  rac_assemble_input clears the B records it is about to use with a
// cudaMemsetAsync on `stream`, which is what makes the array reusable although it still holds the previous call's sums.
int rac_assemble_input_static(const float* d_init_hwc, const float* d_gt_hwc, const float* d_depth, int B, int W, int H,
                              int use_norm_depth, float* d_x_nchw, rac_stream_t stream_) {
    RAC_REQUIRE(B > 0, "rac_assemble_input_static: bad batch size B=%d", B);
    RAC_REQUIRE(B <= RAC_ASSEMBLE_STATIC_MAX_B,
                "rac_assemble_input_static: B=%d exceeds the built-in scratch (RAC_ASSEMBLE_STATIC_MAX_B = %d images); "
                "split the batch or use rac_assemble_input with scratch of your own", B, RAC_ASSEMBLE_STATIC_MAX_B);
    void* scratch = nullptr;
    const int rc = assemble_static_scratch(&scratch);
    if (rc != RAC_OK) return rc;
    return rac_assemble_input(d_init_hwc, d_gt_hwc, d_depth, B, W, H, use_norm_depth, d_x_nchw, scratch,
                              sizeof(g_assemble_static), stream_);
}

}  // extern "C"
