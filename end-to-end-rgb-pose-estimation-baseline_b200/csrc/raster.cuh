// raster.cuh — data layout shared by the render kernels (setup -> band raster -> finalize).
//
// Pipeline spec ("RS", restated from SURVEY.md §8a R1-R9; the CPU oracle in oracle/raster_ref.c
// implements the same spec independently):
//   RS-V  vertex:  pc = R p + t (fma chain x,y,z order), iz = 1/zc, u = fma(fx, xc*iz, cx),
//         v = fma(fy, yc*iz, cy); snapped to 1/256 px: X = rint(256 u), Y = rint(256 v).
//         Near plane: GL clips a triangle at z = znear; here every SAMPLE with d < 0 (z < znear) is rejected by the depth
//         test (RS-Z), which is the same set of samples, as long as the triangle can be projected at all: a triangle
//         with any vertex at zc < znear / 64 (0.8 mm for pyrender's 0.05 m) or with |u|,|v| > 16384 px is dropped whole.
//   RS-T  triangle: area2 = (X1-X0)(Y2-Y0) - (X2-X0)(Y1-Y0) in int64; front-facing iff
//         area2 < 0 (GL CCW after the v-flip, R4); v1<->v2 are swapped so the interior is E>=0.
//   RS-C  coverage: 4 samples per pixel at subpixel (96,32),(224,96),(32,160),(160,224)/256,
//         exact int64 edge functions, top-left fill rule.
//   RS-Z  depth: 1/z is affine in screen space: iz(P) = fma(gy, Py-Y0, fma(gx, Px-X0, iz0));
//         d = fma(-fn/(f-n), iz, f/(f-n)); d24 = rint(d * (2^24-1)); GL_LESS against 0xFFFFFF;
//         ties broken by the lower triangle index (GL's in-order rasterisation).
//   RS-S  shading once per (pixel, triangle) at the pixel centre, pyrender mesh.frag
//         metallic-roughness BRDF with a head-light spot (R7), gamma 1/2.2, RGBA8 round-to-nearest.
//   RS-R  resolve: colour = (sum of the 4 samples' RGBA8 + 2) >> 2 (background 0); depth = sample 0.
//   RS-D  depth readback (R6): d = d24/(2^24-1); zn = 2d-1; depth = 2nf / ((f+n) - zn (f-n)) in f32.
//   RS-N  per-image max-normalisation of the 8-bit colour (R9).
#pragma once
#include "rac_common.cuh"

namespace rac {

constexpr int kSubpix = 256;          // 8 subpixel bits
constexpr float kGuardPx = 16384.0f;  // guard band in pixels
constexpr float kNearRelax = 1.0f / 64.0f;   // vertices are projected down to znear / 64; the near plane clips per sample
constexpr unsigned kDepthMax = 0xFFFFFFu;
constexpr unsigned long long kKeyEmpty = ~0ull;
constexpr int kTileW = 32;            // raster tile in pixels: 32 x 16 (visibility buffer: 512 px * 4 samples * 8 B = 16 KB);
constexpr int kTileH = 16;            // 32 px wide keeps every output row segment one 128-B line
constexpr int kSetupThreads = 256;    // triangles per setup CTA (= refs per descriptor, u8 local ids)
constexpr int kMaxBinTiles = 2048;    // images with more tiles (> 2 Mpx) skip binning: every triangle is "large"

// per-image accumulators, zeroed at the start of every call
struct ImgStats {
    unsigned nlarge;     // large triangles of the image (all setup CTAs); 0 lets the bins skip the large-list scan
    unsigned qmax;       // max resolved 8-bit colour value (R9)
    unsigned dcount;     // #pixels with depth > 0.01      (normalize_depth mask)
    unsigned dmax_bits;  // float bits of max depth (depth >= 0 so uint order == float order)
    double dsum, dsumsq; // over depth > 0.01
    unsigned ndone;      // setup CTAs of this image that have finished (the last one publishes the image's bins)
    unsigned pad0;
    double pad1[3];
};
static_assert(sizeof(ImgStats) == 64, "ImgStats must be 64 B");

// One record per (image, triangle) that survived setup: 3 x 16 B.
//  q0 = {X0, Y0, X1, Y1}   q1 = {X2, Y2, iz0, gx}   q2 = {gy, nx, ny, nz}
struct TriRec {
    int4 q0, q1, q2;
};

// Tile grid of one image.  binned == 0 (more than kMaxBinTiles tiles): every survivor goes to the large list.
struct TileGrid {
    int ntx, nty, nt;
    int binned;
    int ncta_max;        // ceil(maxT / kSetupThreads): stride of the descriptor / ref / large-list tables
};

// Device-side work-list control.  The non-empty bins of the batch are kept in kBuckets lists by load class
// (bucket 0 = heaviest): the setup CTA that finishes an image last appends that image's bins to the lists, and the
// persistent raster CTAs consume the concatenation bucket 0, 1, ... — heaviest bins first (LPT), which keeps the
// tail of raster_kernel short when there are only a few bins per CTA, without a sorting pass between the kernels.
constexpr int kBuckets = 16;
constexpr int kBucketShift = 5;     // load class = min(load >> 5, 15)
struct WorkCtl {
    unsigned pad_a;
    unsigned next;       // next raster item
    unsigned pad[62];    // (diagnostic builds accumulate phase cycle counts here)
    unsigned bucket_cnt[kBuckets];
    unsigned pad2[64 - kBuckets];
};
static_assert(sizeof(WorkCtl) == 512, "WorkCtl must be 512 B");

struct RenderWs {
    ImgStats* stats;
    int* mesh_ids;
    unsigned* large_cnt;   // [B][ncta_max] large triangles found by each setup CTA
    unsigned* large_tri;   // [B][maxT] their ids: slice [cta * 256, cta * 256 + large_cnt) per setup CTA
    int4* recs;            // [B][maxT][3]
    unsigned* desc;        // [B][nt][ncta_max]  (offset << 16) | count, one per (bin, setup CTA)
    unsigned char* refs;   // [B][ncta_max][1024] local triangle ids grouped by bin
    WorkCtl* ctl;          // work-list counters (zeroed with stats and loads by one memset)
    unsigned* loads;       // [B][nt] load estimate of the bin (0 = empty bin)
    unsigned* buckets;     // [kBuckets][B * nt] non-empty bins (image * nt + bin) by load class
    unsigned* rgba;        // [B][H][W] resolved RGBA8 (non-empty bins only)
    float* depth;          // [B][H][W] metric depth (non-empty bins only)
    size_t zero_bytes;     // bytes from `stats` that every call zeroes (stats + ctl + loads)
    size_t total;
};

struct SceneConsts {
    float znear, zfar;
    float c0, c1;            // d = c0 - c1 * iz  (f/(f-n), fn/(f-n))
    float rb_num, rb_sum, rb_diff;  // depth readback constants 2nf, f+n, f-n (f32, numpy weak-scalar casts)
    float las, lao;          // spot-light angle scale / offset
    float intensity, ambient;
};

}  // namespace rac
