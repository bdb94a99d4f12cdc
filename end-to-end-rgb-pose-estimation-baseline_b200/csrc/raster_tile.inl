// raster_tile.inl — raster_kernel: one CTA per (image, 32x16-px bin) work item.  Spec: RS-C, RS-Z, RS-R in raster.cuh.
// Included by raster.cu (inside namespace rac).
//
//   gather    the bin's descriptors (one per setup CTA) -> u8 ref regions -> triangle ids in a shared queue;
//   coverage  batches of <= 254 triangles, thread = triangle: triangles with <= 20 bbox pixels are staged (int32 edge
//             equations relative to the clipped bbox) and their (triangle, pixel) items flattened over the warp, 32
//             per step; larger ones are walked by a whole warp each.  Exact edge functions, top-left rule, 4 samples;
//   depth     24-bit depth and the triangle's batch-local id form one 32-bit key (d24 << 8 | id): the z-test is ONE
//             native shared-memory atomic (ATOMS.MIN.U32), no 64-bit CAS loop.  Ids follow the triangle order, so
//             ties go to the lower triangle index (GL's in-order rasterisation);
//   resolve   per pixel the distinct winners among the 4 samples are listed, the (pixel, triangle) items are shaded
//             256 per step (flattened again), colours averaged, RGBA8 + f32 depth stored as 128-B row segments, and
//             the per-image max / depth statistics reduced -> one atomic per accumulator per bin.
// A bin with more than 254 triangles commits between batches: the current winners are shaded, their colours parked in
// the shared-memory colour buffer of the resolve (at the samples they won) and their ids cleared (id 0 = committed: a later triangle at the same depth loses, as
// in-order rasterisation would have it).  The warp-autonomous alternative (one warp per 16x8 sub-tile, no barriers)
// was built and measured slower: see profiles/README.md.

// Optional phase timeline (build with -DRAC_PHASE_TIMING): thread 0 of every CTA accumulates clock64() deltas per
// phase into WorkCtl::pad (read back from the workspace by profiles/phase_timing.py).  Off in the product build.
#ifdef RAC_PHASE_TIMING
#define RAC_PHASE(idx)                                                                                          \
    do {                                                                                                        \
        if (threadIdx.x == 0) {                                                                                 \
            const long long _t = clock64();                                                                     \
            if (g_phase_ctl) atomicAdd(reinterpret_cast<unsigned long long*>(g_phase_ctl->pad) + (idx), (unsigned long long)(_t - g_phase_t)); \
            g_phase_t = _t;                                                                                     \
        }                                                                                                       \
    } while (0)
#define RAC_PHASE_ARGS , WorkCtl* g_phase_ctl, long long& g_phase_t
#define RAC_PHASE_PASS , g_phase_ctl, g_phase_t
#else
#define RAC_PHASE(idx) do { } while (0)
#define RAC_PHASE_ARGS
#define RAC_PHASE_PASS
#endif

constexpr int kRasterThreads = 256;
constexpr int kRasterWarpsN = kRasterThreads / 32;
constexpr int kQueueCap = 1024;       // triangle ids gathered per round
constexpr int kBatch = 254;           // triangles per coverage batch = local ids 1..254 (0 = committed, 0xFF = the empty key)
#ifndef RAC_SMALL_PIX
#define RAC_SMALL_PIX 20
#endif
constexpr int kSmallPix = RAC_SMALL_PIX;   // bbox∩bin pixels handled as flattened items; larger -> one warp per triangle
constexpr int kTilePix = kTileW * kTileH;
constexpr int kTriStride = 20;        // words per staged triangle (16 used; 80 B keeps LDS.128 aligned and conflict-free)
constexpr unsigned kKeyEmpty32 = 0xFFFFFFFFu;

// dynamic shared memory of one CTA
struct __align__(16) RasterSmem {
    unsigned z[4 * kTilePix];                      //  8 KB keys, sample-major: z[s * 512 + p]
    int queue[kQueueCap];                          //  4 KB gathered triangle ids
    int medq[kRasterThreads];                      //  1 KB batch-local ids of medium triangles
    int tri[kRasterThreads * kTriStride];          // 20 KB staged triangles by local id - 1: edge equations (coverage) + plane, normal (shading)
    unsigned short items[kRasterThreads * kSmallPix];       // 10 KB flattened (triangle, pixel) items of the batch; resolve: (pixel, sample) items
    unsigned slots[kBatch + 2];                    //  1 KB local id - 1 -> triangle id
    unsigned cols[4 * kTilePix];                   //  8 KB resolve: colour of each pixel's representative samples
};
constexpr size_t kRasterSmemBytes = sizeof(RasterSmem);

// bbox (in pixels) of the triangle clipped to the bin [x_lo, x_hi] x [y_lo, y_hi]
__device__ __forceinline__ void tri_bbox(const TriGeom& g, int x_lo, int x_hi, int y_lo, int y_hi,
                                         int& bx0, int& bx1, int& by0, int& by1) {
    const int xmin = min(g.X0, min(g.X1, g.X2)), xmax = max(g.X0, max(g.X1, g.X2));
    const int ymin = min(g.Y0, min(g.Y1, g.Y2)), ymax = max(g.Y0, max(g.Y1, g.Y2));
    bx0 = max(ceil_div256(xmin - 224), x_lo);
    bx1 = min(floor_div256(xmax - 32), x_hi);
    by0 = max(ceil_div256(ymin - 224), y_lo);
    by1 = min(floor_div256(ymax - 32), y_hi);
}

// RS-Z for one covered sample -> key, or kKeyEmpty32 when the depth is outside [0, 1)
__device__ __forceinline__ unsigned depth_key(float iz, float c0, float c1, unsigned local) {
    const float d = __fmaf_rn(-c1, iz, c0);
    if (d >= 0.0f && d <= 1.0f) {
        const unsigned d24 = (unsigned)__float2int_rn(__fmul_rn(d, 16777215.0f));
        if (d24 < kDepthMax) return (d24 << 8) | local;
    }
    return kKeyEmpty32;
}

// RS-C + RS-Z for one triangle over its bbox∩bin, one warp cooperating (lane = pixel).
// I = int for binned triangles (extent <= 64 px: every product < 2^29), long long for large ones.
template <typename I>
__device__ __forceinline__ void raster_tri_warp(const TriGeom& g, unsigned local, int bx0, int bx1, int by0, int by1,
                                                int lane, unsigned* z, int x_lo, int y_lo, float c0, float c1) {
    // edges e0: v0->v1, e1: v1->v2, e2: v2->v0;  E(P) = dx (Py-ay) - dy (Px-ax), interior E >= 0
    const I dx0 = g.X1 - g.X0, dy0 = g.Y1 - g.Y0;
    const I dx1 = g.X2 - g.X1, dy1 = g.Y2 - g.Y1;
    const I dx2 = g.X0 - g.X2, dy2 = g.Y0 - g.Y2;
    // top-left rule: samples exactly on an edge belong to the triangle only for top/left edges
    const I b0 = (dy0 < 0 || (dy0 == 0 && dx0 > 0)) ? 0 : -1;
    const I b1 = (dy1 < 0 || (dy1 == 0 && dx1 > 0)) ? 0 : -1;
    const I b2 = (dy2 < 0 || (dy2 == 0 && dx2 > 0)) ? 0 : -1;
    const int w = bx1 - bx0 + 1;
    const int n = w * (by1 - by0 + 1);
    constexpr int SX[4] = {96, 224, 32, 160};
    constexpr int SY[4] = {32, 96, 160, 224};
    for (int i = lane; i < n; i += 32) {
        const int ry = i / w;
        const int px = bx0 + (i - ry * w), py = by0 + ry;
        const int ox = px * kSubpix, oy = py * kSubpix;
        const I E0 = dx0 * (I)(oy - g.Y0) - dy0 * (I)(ox - g.X0) + b0;
        const I E1 = dx1 * (I)(oy - g.Y1) - dy1 * (I)(ox - g.X1) + b1;
        const I E2 = dx2 * (I)(oy - g.Y2) - dy2 * (I)(ox - g.X2) + b2;
        const int pidx = (py - y_lo) * kTileW + (px - x_lo);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const I e0 = E0 + dx0 * (I)SY[s] - dy0 * (I)SX[s];
            const I e1 = E1 + dx1 * (I)SY[s] - dy1 * (I)SX[s];
            const I e2 = E2 + dx2 * (I)SY[s] - dy2 * (I)SX[s];
            if ((e0 | e1 | e2) >= 0) {
                const float fxr = (float)(ox + SX[s] - g.X0), fyr = (float)(oy + SY[s] - g.Y0);
                const float iz = __fmaf_rn(g.gy, fyr, __fmaf_rn(g.gx, fxr, g.iz0));
                const unsigned key = depth_key(iz, c0, c1, local);
                if (key != kKeyEmpty32) atomicMin(&z[s * kTilePix + pidx], key);
            }
        }
    }
}

// per-bin constants every phase needs
struct BinCtx {
    int x_lo, y_lo, x_hi, y_hi;      // inclusive pixel bounds (clipped to the image)
    float c0, c1;
    const int4* img_recs;
};

// ---------------------------------------------------------------------------------------------
// coverage of one batch: ids[0..n), n <= kBatch, local ids first_local .. first_local + n - 1 (thread = triangle)
// ---------------------------------------------------------------------------------------------
// Binned triangles (int32-exact) are flattened so that lanes stay busy: each thread stages its triangle (edge
// equations relative to its clipped bbox origin, depth plane, id) in shared memory and lists the bbox pixels it must
// visit; the warp then walks the concatenated (triangle, pixel) items of its 32 triangles, 32 per step, so a 1-pixel
// triangle and a 16-pixel triangle in the same warp cost 17 items, not 2 x 16.  Triangles with more than kSmallPix
// bbox pixels (and every LARGE one: int64 edge functions) take the warp-per-triangle path.  Called by all threads.
template <bool LARGE>
__device__ __forceinline__ void process_batch(RasterSmem& sm, int* med_count, unsigned* s_wsum, const BinCtx& bc, const int* ids,
                                              int n, unsigned first_local RAC_PHASE_ARGS) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int SX[4] = {96, 224, 32, 160};
    constexpr int SY[4] = {32, 96, 160, 224};
    int nit = 0, w = 0, h = 0;
    const int slot_id = (int)first_local - 1 + tid;       // = local id - 1 (< 254)
    if (tid < n) {
        const unsigned tri = (unsigned)ids[tid];
        const unsigned local = first_local + (unsigned)tid;
        sm.slots[local - 1u] = tri;
        const int4* r = bc.img_recs + 3 * (size_t)tri;
        const int4 q0 = r[0], q1 = r[1], q2 = r[2];
        const TriGeom g = unpack_geom(q0, q1, q2);
        int bx0, bx1, by0, by1;
        tri_bbox(g, bc.x_lo, bc.x_hi, bc.y_lo, bc.y_hi, bx0, bx1, by0, by1);
        if (bx0 <= bx1 && by0 <= by1) {
            w = bx1 - bx0 + 1; h = by1 - by0 + 1;
            // the slot of this local id: shading data always (it lives until the next commit / the final resolve),
            // edge equations relative to the clipped bbox origin for the flattened coverage pass
            int4* slot = reinterpret_cast<int4*>(sm.tri + (local - 1u) * kTriStride);
            slot[3] = make_int4(__float_as_int(g.gy), __float_as_int(g.iz0), g.X0, g.Y0);
            slot[4] = make_int4(q2.y, q2.z, q2.w, __float_as_int(g.gx));
            if (!LARGE && w * h <= kSmallPix && w <= 16 && h <= 16) {      // (dx, dy are 4-bit fields of an item)
                nit = w * h;
                const int dx0 = g.X1 - g.X0, dy0 = g.Y1 - g.Y0;
                const int dx1 = g.X2 - g.X1, dy1 = g.Y2 - g.Y1;
                const int dx2 = g.X0 - g.X2, dy2 = g.Y0 - g.Y2;
                const int b0 = (dy0 < 0 || (dy0 == 0 && dx0 > 0)) ? 0 : -1;
                const int b1 = (dy1 < 0 || (dy1 == 0 && dx1 > 0)) ? 0 : -1;
                const int b2 = (dy2 < 0 || (dy2 == 0 && dx2 > 0)) ? 0 : -1;
                const int ox = bx0 * kSubpix, oy = by0 * kSubpix;
                const int C0 = dx0 * (oy - g.Y0) - dy0 * (ox - g.X0) + b0;
                const int C1 = dx1 * (oy - g.Y1) - dy1 * (ox - g.X1) + b1;
                const int C2 = dx2 * (oy - g.Y2) - dy2 * (ox - g.X2) + b2;
                slot[0] = make_int4(-dy0, dx0, C0, -dy1);
                slot[1] = make_int4(dx1, C1, -dy2, dx2);
                slot[2] = make_int4(C2, ox, oy, (by0 - bc.y_lo) * kTileW + (bx0 - bc.x_lo));
            } else {
                sm.medq[atomicAdd(med_count, 1)] = tid;
            }
        }
    }
    if (!LARGE) {
        // CTA-wide item list: the (triangle, pixel) items of the whole batch are concatenated and walked 256 per step,
        // so no warp waits at the barrier below for a neighbour that happened to draw the 16-pixel triangles
        int incl = nit;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_wsum[warp] = (unsigned)incl;
        __syncthreads();
        int wbase = 0, total = 0;
#pragma unroll
        for (int w2 = 0; w2 < kRasterWarpsN; ++w2) {
            const int v = (int)s_wsum[w2];
            if (w2 < warp) wbase += v;
            total += v;
        }
        if (nit) {
            int idx = wbase + incl - nit;
            for (int dy = 0; dy < h; ++dy)
                for (int dx = 0; dx < w; ++dx) sm.items[idx++] = (unsigned short)(slot_id | (dx << 8) | (dy << 12));
        }
        __syncthreads();
        RAC_PHASE(13);
        const float c0 = bc.c0, c1 = bc.c1;
        for (int j = tid; j < total; j += kRasterThreads) {
            const int code = sm.items[j];
            const int dxp = (code >> 8) & 15, dyp = code >> 12;
            const int4* slot = reinterpret_cast<const int4*>(sm.tri + (code & 255) * kTriStride);
            const int4 a = slot[0], b = slot[1], c = slot[2], d = slot[3];
            const int sx = dxp * kSubpix, sy = dyp * kSubpix;
            const float gx = __int_as_float(slot[4].w), gy = __int_as_float(d.x), iz0 = __int_as_float(d.y);
            const int rx = c.y - d.z, ry = c.z - d.w;       // bbox origin relative to vertex 0
            const int pidx = c.w + dyp * kTileW + dxp;
            const unsigned local = (unsigned)(code & 255) + 1u;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int ux = sx + SX[s], uy = sy + SY[s];
                const int e0 = a.z + a.x * ux + a.y * uy;
                const int e1 = b.y + a.w * ux + b.x * uy;
                const int e2 = c.x + b.z * ux + b.w * uy;
                if ((e0 | e1 | e2) >= 0) {
                    const float fxr = (float)(rx + ux), fyr = (float)(ry + uy);
                    const float iz = __fmaf_rn(gy, fyr, __fmaf_rn(gx, fxr, iz0));
                    const unsigned key = depth_key(iz, c0, c1, local);
                    if (key != kKeyEmpty32) atomicMin(&sm.z[s * kTilePix + pidx], key);
                }
            }
        }
    }
    RAC_PHASE(2);
    __syncthreads();
    RAC_PHASE(3);
    const int mn = *med_count;
    for (int k = warp; k < mn; k += kRasterWarpsN) {
        const int t = sm.medq[k];
        const int4* r = bc.img_recs + 3 * (size_t)(unsigned)ids[t];
        const TriGeom g = unpack_geom(r[0], r[1], r[2]);
        int bx0, bx1, by0, by1;
        tri_bbox(g, bc.x_lo, bc.x_hi, bc.y_lo, bc.y_hi, bx0, bx1, by0, by1);
        raster_tri_warp<typename std::conditional<LARGE, long long, int>::type>(
            g, first_local + (unsigned)t, bx0, bx1, by0, by1, lane, sm.z, bc.x_lo, bc.y_lo, bc.c0, bc.c1);
    }
    __syncthreads();
    RAC_PHASE(4);
    // (no barrier after the reset: consecutive batches alternate between two counters, and this one is next bumped
    // two batches — several barriers — from here)
    if (tid == 0) *med_count = 0;
}

// ---------------------------------------------------------------------------------------------
// resolve / commit
// ---------------------------------------------------------------------------------------------
struct TileAcc {
    unsigned qmax, dcount;
    float dmax;
    double dsum, dsumsq;
};

// R2 of resolve_bin: shade the listed (pixel, representative sample) items, 256 per step.
template <bool FINAL, bool PARTS>
__device__ __forceinline__ void shade_items(RasterSmem& sm, const BinCtx& bc, const ShadeCtx* s_sc, int nitems) {
    const int tid = threadIdx.x;
    const unsigned short* sitems = sm.items;
    unsigned* cols = sm.cols;
    const ShadeCtx& sc = *s_sc;     // built once per bin by the last thread, long before this point; read in place
                                    // (shared memory): keeps ~28 values out of the registers of the shading loop
    for (int it = tid; it < nitems; it += kRasterThreads) {
        const int code = sitems[it];
        const int p = code & 1023, s = code >> 10;
        const unsigned key = sm.z[s * kTilePix + p];
        const unsigned id1 = (key & 0xFFu) - 1u;
        const unsigned tri = sm.slots[id1];
        const int4* slot = reinterpret_cast<const int4*>(sm.tri + id1 * kTriStride);
        const int4 d = slot[3], e = slot[4];
        ShadeTri st;
        st.X0 = d.z; st.Y0 = d.w; st.gy = __int_as_float(d.x); st.iz0 = __int_as_float(d.y);
        st.nx = __int_as_float(e.x); st.ny = __int_as_float(e.y); st.nz = __int_as_float(e.z); st.gx = __int_as_float(e.w);
        const unsigned col = shade_t<PARTS>(sc, st, bc.img_recs + 3 * (size_t)tri, tri, bc.x_lo + (p & (kTileW - 1)), bc.y_lo + p / kTileW);
        if (FINAL) {
            cols[s * kTilePix + p] = col;      // R3 finds it through the representative index
        } else {
            // commit: every sample of this pixel won by the same triangle gets the colour; the ids are cleared
#pragma unroll
            for (int s2 = 0; s2 < 4; ++s2) {
                const unsigned k2 = sm.z[s2 * kTilePix + p];
                if (s2 >= s && k2 != kKeyEmpty32 && ((k2 ^ key) & 0xFFu) == 0u) {
                    cols[s2 * kTilePix + p] = col;     // stays there: only the final resolve writes cols again, and only
                                                       // at representative samples of live ids
                    sm.z[s2 * kTilePix + p] = k2 & ~0xFFu;
                }
            }
        }
    }
}
template <bool FINAL>
__device__ __noinline__ void shade_items_parts(RasterSmem& sm, const BinCtx& bc, const ShadeCtx* s_sc, int nitems) {
    shade_items<FINAL, true>(sm, bc, s_sc, nitems);
}

// Flattened shading.  R1: every pixel finds the distinct uncommitted winners among its 4 samples and lists one
// (pixel, representative sample) item per distinct triangle; R2: the items are shaded 256 at a time (no lane idles on
// background or waits for a neighbour's 4-triangle edge pixel); R3 (FINAL): every pixel averages its samples' colours
// through the representatives / the committed colours and stores RGBA8 + depth.
// FINAL = false (commit): R2 parks the colours in `cols`, at the samples the triangle won, and clears the ids; no R3.
// (`cols` is written again only by the final R2, and only at representative samples of live ids, so a parked colour
// survives exactly as long as its sample does.)
// Called by all threads; q_count is a zeroed shared counter (left zeroed).
template <bool FINAL>
__device__ __forceinline__ void resolve_bin(RasterSmem& sm, int* q_count, const BinCtx& bc,
                                            const ShadeCtx* s_sc,
                                            const SceneConsts& S, unsigned* __restrict__ rgba_out,
                                            float* __restrict__ depth_out, size_t img_pix, int W, TileAcc& acc, int need_dstats
                                            RAC_PHASE_ARGS) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int tw = bc.x_hi - bc.x_lo + 1, th = bc.y_hi - bc.y_lo + 1;
    unsigned short* sitems = sm.items;   // [4 * kTilePix] of the 4096: free between batches (the queue is NOT: a commit runs mid-queue)
    unsigned* cols = sm.cols;            // [4 * kTilePix] colour of each representative sample
    constexpr int kPer = kTilePix / kRasterThreads;
    // R1.  Nothing is kept in registers across the shading pass: R3 re-derives the representatives from the keys.
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
        const int p = tid + i * kRasterThreads;
        const int lx = p & (kTileW - 1), ly = p / kTileW;
        const bool inb = lx < tw && ly < th;
        unsigned k[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) k[s] = inb ? sm.z[s * kTilePix + p] : kKeyEmpty32;
        // a warp owns one 32-pixel row segment: nothing to list when all of it is background (most of a silhouette bin)
        if (!__any_sync(0xffffffffu, (k[0] & k[1] & k[2] & k[3]) != kKeyEmpty32)) continue;
        unsigned fresh = 0;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            if (k[s] == kKeyEmpty32 || (k[s] & 0xFFu) == 0u) continue;
            bool first = true;
#pragma unroll
            for (int j = s - 1; j >= 0; --j)
                if (((k[j] ^ k[s]) & 0xFFu) == 0u) first = false;      // (ids are 1..254: 0 = committed, 0xFF = empty key)
            if (first) fresh |= 1u << s;
        }
        // append this pixel's items (warp-aggregated)
        const int c = __popc(fresh);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        int base = 0;
        if (lane == 31 && incl) base = atomicAdd(q_count, incl);
        base = __shfl_sync(0xffffffffu, base, 31) + incl - c;
#pragma unroll
        for (int s = 0; s < 4; ++s)
            if (fresh & (1u << s)) sitems[base++] = (unsigned short)(p | (s << 10));
    }
    RAC_PHASE(6);
    __syncthreads();
    RAC_PHASE(7);
    const int nitems = *q_count;
    if (nitems > 0) {       // CTA-uniform
        // multi-part meshes (GLB scenes: material / texture per face) take an out-of-line copy of the loop, so the
        // single-material path keeps its register allocation
        if (s_sc->parts) shade_items_parts<FINAL>(sm, bc, s_sc, nitems);
        else shade_items<FINAL, false>(sm, bc, s_sc, nitems);
    }
    RAC_PHASE(8);
    __syncthreads();
    RAC_PHASE(9);
    if (tid == 0) *q_count = 0;
    if (!FINAL) { __syncthreads(); return; }
#pragma unroll
    for (int i = 0; i < kPer; ++i) {
        const int p = tid + i * kRasterThreads;
        const int lx = p & (kTileW - 1), ly = p / kTileW;
        const bool inb = lx < tw && ly < th;
        unsigned k[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) k[s] = inb ? sm.z[s * kTilePix + p] : kKeyEmpty32;
        if (!__any_sync(0xffffffffu, (k[0] & k[1] & k[2] & k[3]) != kKeyEmpty32)) {     // background row segment
            if (inb) {
                const size_t o = img_pix + (size_t)(bc.y_lo + ly) * W + (bc.x_lo + lx);
                rgba_out[o] = 0u;
                if (depth_out) depth_out[o] = 0.0f;
            }
            continue;
        }
        if (!inb) continue;
        unsigned sum_r = 0, sum_g = 0, sum_b = 0, sum_a = 0;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            if (k[s] == kKeyEmpty32) continue;
            unsigned col;
            if ((k[s] & 0xFFu) == 0u) {
                col = cols[s * kTilePix + p];          // committed earlier: parked in place by the commit
            } else {
                int rep = s;
#pragma unroll
                for (int j = s - 1; j >= 0; --j)
                    if (((k[j] ^ k[s]) & 0xFFu) == 0u) rep = j;
                col = cols[rep * kTilePix + p];
            }
            sum_r += col & 0xFFu; sum_g += (col >> 8) & 0xFFu;
            sum_b += (col >> 16) & 0xFFu; sum_a += col >> 24;
        }
        const unsigned r8 = (sum_r + 2u) >> 2, g8 = (sum_g + 2u) >> 2, b8 = (sum_b + 2u) >> 2, a8 = (sum_a + 2u) >> 2;
        const size_t o = img_pix + (size_t)(bc.y_lo + ly) * W + (bc.x_lo + lx);
        rgba_out[o] = r8 | (g8 << 8) | (b8 << 16) | (a8 << 24);
        acc.qmax = max(acc.qmax, max(r8, max(g8, b8)));
        if (depth_out) {        // kernel-uniform: nobody reads the metric depth of a 6-channel concat without depth output
            float depth = 0.0f;
            if (k[0] != kKeyEmpty32) depth = depth_readback(k[0] >> 8, S);
            depth_out[o] = depth;
            if (need_dstats) {  // normalize_depth statistics: only when the caller asked for the normalised depth plane
                acc.dmax = fmaxf(acc.dmax, depth);
                if (depth > 0.01f) { acc.dcount++; acc.dsum += (double)depth; acc.dsumsq += (double)depth * (double)depth; }
            }
        }
    }
}

// commit (rare: more than 254 triangles in a bin) — out of line, so that the hot path carries one copy of the resolve
__device__ __noinline__ void commit_bin(RasterSmem& sm, int* q_count, const BinCtx& bc,
                                        const ShadeCtx* s_sc, const SceneConsts& S RAC_PHASE_ARGS) {
    TileAcc dummy;
    resolve_bin<false>(sm, q_count, bc, s_sc, S, nullptr, nullptr, 0, 0, dummy, 0 RAC_PHASE_PASS);
}

// ---------------------------------------------------------------------------------------------
// one bin, start to finish
// ---------------------------------------------------------------------------------------------
// s_ctl: [0] resolve item count, [1], [2] medium-triangle counts of alternating batches (all zero on entry and on exit);
// s_scan: [4 * kRasterThreads + 8] descriptor counts / scan / offsets / large counts / warp sums.
__device__ __forceinline__ void raster_tile(RasterSmem& sm, int* s_ctl, unsigned* s_scan, ShadeCtx* s_sc,
                                            const MeshDev* __restrict__ table, const int* __restrict__ mesh_ids,
                                            int uniform_mesh, const float* __restrict__ T_CO, const float* __restrict__ K,
                                            int W, int H, int maxT, const TileGrid& tg, const SceneConsts& S,
                                            ImgStats* __restrict__ stats, const unsigned* __restrict__ large_cnt,
                                            const unsigned* __restrict__ large_tri, const int4* __restrict__ recs,
                                            const unsigned* __restrict__ desc, const unsigned char* __restrict__ refs,
                                            unsigned* __restrict__ rgba_out, float* __restrict__ depth_out,
                                            int b, int tile, unsigned r_next, unsigned n_work, const unsigned* s_prefix,
                                            const unsigned* __restrict__ buckets, size_t bucket_cap, unsigned* s_next,
                                            int need_dstats RAC_PHASE_ARGS) {
    int* q_count = s_ctl, *med_count = s_ctl + 1;
    int batch_no = 0;          // CTA-uniform: selects the medium-triangle counter of the batch
    unsigned* s_dcnt = s_scan;                      // [kRasterThreads] descriptor counts (large pass: per-CTA counts)
    unsigned* s_dscan = s_scan + kRasterThreads;    // [kRasterThreads] exclusive scan
    unsigned* s_doff = s_scan + 2 * kRasterThreads; // [kRasterThreads] offset into the ref region
    unsigned* s_wsum = s_scan + 3 * kRasterThreads; // [8]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    BinCtx bc;
    bc.x_lo = (tile % tg.ntx) * kTileW; bc.y_lo = (tile / tg.ntx) * kTileH;
    bc.x_hi = min(bc.x_lo + kTileW, W) - 1; bc.y_hi = min(bc.y_lo + kTileH, H) - 1;
    bc.c0 = S.c0; bc.c1 = S.c1;
    bc.img_recs = recs + 3 * (size_t)b * maxT;
    const MeshDev& mesh = table[mesh_ids ? mesh_ids[b] : uniform_mesh];
    const int ncta = (mesh.T + kSetupThreads - 1) / kSetupThreads;
    const size_t img_pix = (size_t)b * W * H;
    // the bin's first descriptor group is requested before anything else: its L2 round trip overlaps the z-buffer
    // initialisation, the shading-context loads of the last thread and the first barrier
    const unsigned* dtile = desc + ((size_t)b * tg.nt + tile) * tg.ncta_max;
    unsigned d_first = 0u;
    if (tg.binned && tid < ncta) d_first = __ldg(dtile + tid);
    {
        uint4* z4 = reinterpret_cast<uint4*>(sm.z);
        const uint4 e = make_uint4(kKeyEmpty32, kKeyEmpty32, kKeyEmpty32, kKeyEmpty32);
#pragma unroll
        for (int i = 0; i < kTilePix / kRasterThreads; ++i) z4[tid + i * kRasterThreads] = e;
    }
    int nslots = 0;           // local ids handed out since the last commit (CTA-uniform)
    if (tid == kRasterThreads - 1) {     // the shading context of this (image, mesh): consumed after several barriers
        const CamPose cam = load_campose(T_CO, K, b);
        *s_sc = make_shade_ctx(mesh, cam, S);
    }
    const unsigned n_large_img = stats[b].nlarge;     // written by setup_kernel (previous launch)
    // coverage of sm.queue[0..qn) in batches of <= kBatch local ids; commits when the ids run out
    auto run_batches = [&](int qn, auto large_tag) {
        constexpr bool LARGE = decltype(large_tag)::value;
        for (int qb = 0; qb < qn; qb += kBatch) {
            const int n = min(kBatch, qn - qb);
            if (nslots + n > kBatch) {
                commit_bin(sm, q_count, bc, s_sc, S RAC_PHASE_PASS);
                nslots = 0;
            }
            process_batch<LARGE>(sm, med_count + (batch_no & 1), s_wsum, bc, sm.queue + qb, n, (unsigned)nslots + 1u RAC_PHASE_PASS);
            ++batch_no;
            nslots += n;
        }
    };
    // (binned images: the descriptor scan below starts with barriers of its own, well before anything reads the
    // z-buffer, the control words or the shading context initialised above)
    if (!tg.binned) __syncthreads();

    // ---- binned triangles: descriptors of this bin, kRasterThreads at a time ----
    if (tg.binned) {
        for (int dbase = 0; dbase < ncta; dbase += kRasterThreads) {
            const int j = dbase + tid;
            const unsigned d = dbase == 0 ? d_first : (j < ncta ? dtile[j] : 0u);
            const unsigned c = d & 0xFFFFu;
            unsigned incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            unsigned wbase = 0, total = 0;
#pragma unroll
            for (int w = 0; w < kRasterWarpsN; ++w) {
                const unsigned v = s_wsum[w];
                if (w < warp) wbase += v;
                total += v;
            }
            s_dcnt[tid] = c; s_dscan[tid] = wbase + incl - c; s_doff[tid] = d >> 16;
            __syncthreads();
            const int nd = min(kRasterThreads, ncta - dbase);
            for (unsigned base = 0; base < total; base += kQueueCap) {
                // expand refs [base, base + kQueueCap) of this descriptor group into the queue: one thread per ref, which
                // finds its descriptor by binary search in the scanned counts — every ref byte of the bin is requested
                // in the same L2 round trip (a warp walking its descriptors one after the other paid one trip each)
                {
                    const unsigned qn = min((unsigned)kQueueCap, total - base);
                    const int top = 1 << (31 - __clz(max(nd - 1, 1)));
                    for (unsigned q = (unsigned)tid; q < qn; q += kRasterThreads) {
                        const unsigned g = base + q;
                        int jj = 0;
                        for (int step = top; step > 0; step >>= 1)
                            if (jj + step < nd && s_dscan[jj + step] <= g) jj += step;     // last descriptor starting at or before g
                        const unsigned char* region = refs + (((size_t)b * tg.ncta_max + dbase + jj) << 10) + s_doff[jj];
                        sm.queue[q] = (dbase + jj) * kSetupThreads + (int)region[g - s_dscan[jj]];
                    }
                }
                __syncthreads();
                RAC_PHASE(1);
                run_batches((int)min((unsigned)kQueueCap, total - base), std::false_type());
            }
        }
    }

    // ---- large triangles: every bin scans the image's large list, per setup CTA and in triangle order ----
    for (int dbase = 0; n_large_img != 0u && dbase < ncta; dbase += kRasterThreads) {
        s_dcnt[tid] = (dbase + tid < ncta) ? large_cnt[(size_t)b * tg.ncta_max + dbase + tid] : 0u;
        __syncthreads();
        const int nd = min(kRasterThreads, ncta - dbase);
        int qn = 0;
        for (int jj = 0; jj < nd; ++jj) {
            const int cj = (int)s_dcnt[jj];
            if (cj == 0) continue;                 // CTA-uniform
            const unsigned* lst = large_tri + (size_t)b * maxT + (size_t)(dbase + jj) * kSetupThreads;
            bool accept = false;
            unsigned tri = 0;
            if (tid < cj) {                        // cj <= kSetupThreads == kRasterThreads
                tri = lst[tid];
                const int4* r = bc.img_recs + 3 * (size_t)tri;
                const int4 q0 = r[0], q1 = r[1];
                const int xmin = min(q0.x, min(q0.z, q1.x)), xmax = max(q0.x, max(q0.z, q1.x));
                const int ymin = min(q0.y, min(q0.w, q1.y)), ymax = max(q0.y, max(q0.w, q1.y));
                accept = ceil_div256(xmin - 224) <= bc.x_hi && floor_div256(xmax - 32) >= bc.x_lo &&
                         ceil_div256(ymin - 224) <= bc.y_hi && floor_div256(ymax - 32) >= bc.y_lo;
            }
            const unsigned am = __ballot_sync(0xffffffffu, accept);
            if (lane == 0) s_wsum[warp] = (unsigned)__popc(am);
            __syncthreads();
            unsigned wbase = 0, total = 0;
#pragma unroll
            for (int w = 0; w < kRasterWarpsN; ++w) {
                const unsigned v = s_wsum[w];
                if (w < warp) wbase += v;
                total += v;
            }
            if (accept) sm.queue[qn + (int)wbase + __popc(am & ((1u << lane) - 1u))] = (int)tri;
            qn += (int)total;
            __syncthreads();
            if (qn > kQueueCap - kRasterThreads) { run_batches(qn, std::true_type()); qn = 0; }
        }
        if (qn > 0) run_batches(qn, std::true_type());
        __syncthreads();
    }

    RAC_PHASE(5);
    // The ticket of the next bin was claimed when this one started, so it has long arrived: park it in shared memory and
    // fetch its work-list entry with an asynchronous 4-byte copy that lands during the resolve (nothing of either is
    // held in registers across the shading loop, and the L2 round trip is off the hand-over path).
    if (tid == 0) {
        s_next[0] = r_next;
        if (r_next < n_work) {
            int k = 0;
#pragma unroll
            for (int step = kBuckets / 2; step > 0; step >>= 1)
                if (s_prefix[k + step] <= r_next) k += step;
            const unsigned* src = buckets + (size_t)k * bucket_cap + (r_next - s_prefix[k]);
            const unsigned dst = (unsigned)__cvta_generic_to_shared(s_next + 1);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst), "l"(src) : "memory");
        }
    }
    // ---- resolve + shade + write, and per-image accumulators ----
    TileAcc acc;
    acc.qmax = 0; acc.dcount = 0; acc.dmax = 0.0f; acc.dsum = 0.0; acc.dsumsq = 0.0;
    resolve_bin<true>(sm, q_count, bc, s_sc, S, rgba_out, depth_out, img_pix, W, acc, need_dstats RAC_PHASE_PASS);
    RAC_PHASE(10);
    // block reduction -> one atomic per accumulator per CTA
    __shared__ double red_d[2][kRasterWarpsN];
    __shared__ unsigned red_u[3][kRasterWarpsN];
    acc.qmax = warp_max(acc.qmax);
    unsigned dmax_bits = 0u;
    if (need_dstats) {      // kernel-uniform
        acc.dcount = (unsigned)__reduce_add_sync(0xffffffffu, acc.dcount);
        dmax_bits = warp_max(__float_as_uint(acc.dmax));
        acc.dsum = warp_sum(acc.dsum); acc.dsumsq = warp_sum(acc.dsumsq);
    }
    if (lane == 0) {
        red_u[0][warp] = acc.qmax; red_u[1][warp] = acc.dcount; red_u[2][warp] = dmax_bits;
        red_d[0][warp] = acc.dsum; red_d[1][warp] = acc.dsumsq;
    }
    __syncthreads();
    if (tid == 0) {
        unsigned qm = 0, dc = 0, dm = 0;
        double a = 0, c = 0;
        for (int w = 0; w < kRasterWarpsN; ++w) {
            qm = max(qm, red_u[0][w]); dc += red_u[1][w]; dm = max(dm, red_u[2][w]);
            a += red_d[0][w]; c += red_d[1][w];
        }
        ImgStats* st = stats + b;
        if (qm) atomicMax(&st->qmax, qm);
        if (dc) { atomicAdd(&st->dcount, dc); atomicAdd(&st->dsum, a); atomicAdd(&st->dsumsq, c); }
        if (dm) atomicMax(&st->dmax_bits, dm);
    }
}

// Persistent CTAs (SMs x resident CTAs) pop non-empty bins from the bucket lists setup_kernel built (heaviest first).  The claim for
// the NEXT item is issued before the current bin is rasterised, so the atomic's round trip to L2 hides behind the bin.
// (kRasterRegs = 56: 4 CTAs/SM leave 8 K registers, one resident gt_norm CTA of 256 threads x 32 registers per SM.)
constexpr int kRasterRegs = 56;
__global__ void __maxnreg__(kRasterRegs)
raster_kernel(const MeshDev* __restrict__ table, const int* __restrict__ mesh_ids, int uniform_mesh,
              const float* __restrict__ T_CO, const float* __restrict__ K, int W, int H, int maxT, TileGrid tg,
              SceneConsts S, ImgStats* __restrict__ stats, const unsigned* __restrict__ large_cnt,
              const unsigned* __restrict__ large_tri, const int4* __restrict__ recs, const unsigned* __restrict__ desc,
              const unsigned char* __restrict__ refs, WorkCtl* __restrict__ ctl, const unsigned* __restrict__ buckets,
              size_t bucket_cap, unsigned* __restrict__ rgba_out, float* __restrict__ depth_out,
              int need_dstats) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RasterSmem& sm = *reinterpret_cast<RasterSmem*>(smem_raw);
    __shared__ int s_ctl[3];
    __shared__ unsigned s_next[2];     // ticket and work-list entry of the CTA's next bin
    __shared__ unsigned s_scan[3 * kRasterThreads + kRasterWarpsN];
    __shared__ ShadeCtx s_sc;
    __shared__ unsigned s_prefix[kBuckets + 1];
    unsigned first_ticket = 0;                       // requested before the bucket counts: the two L2 round trips overlap
    if (threadIdx.x == 0) first_ticket = atomicAdd(&ctl->next, 1u);
    load_bucket_prefix(ctl, s_prefix);               // final: written by setup_kernel (previous launch)
#ifdef RAC_PHASE_TIMING
    WorkCtl* g_phase_ctl = ctl;
    long long g_phase_t = clock64();
#endif
    RAC_TL(2);
#ifdef RAC_PHASE_TIMING
    if (threadIdx.x == 0) {      // sum of CTA start times (ns since the epoch in g_tl[15], set by the host to ~now)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicAdd(&g_tl[9], t - g_tl[15]);
    }
#endif
    // Every ticket, the first included, comes from the counter: a CTA that becomes resident late (the step graph runs
    // the pose / loss kernels and the gt kernel next to this one) must not own one of the heaviest bins — handing the
    // first tickets out by CTA index was measured 2 % slower for exactly that reason.
    if (threadIdx.x == 0) { s_ctl[0] = 0; s_ctl[1] = 0; s_ctl[2] = 0; s_next[0] = first_ticket; }
    __syncthreads();
    const unsigned n_work = s_prefix[kBuckets];
    unsigned r = s_next[0];
    unsigned e = r < n_work ? bucket_entry(s_prefix, buckets, bucket_cap, r) : 0u;   // image * nt + bin, heaviest class first
    while (r < n_work) {
#ifdef RAC_PHASE_TIMING
        const long long t_bin0 = clock64();
#ifdef RAC_PHASE_CLASS      // phase timeline of one load class only
        {
            int k = 0;
            for (int step = kBuckets / 2; step > 0; step >>= 1)
                if (s_prefix[k + step] <= r) k += step;
            g_phase_ctl = (k == RAC_PHASE_CLASS) ? ctl : nullptr;
        }
#endif
#endif
        unsigned r_next = 0;
        if (threadIdx.x == 0) r_next = atomicAdd(&ctl->next, 1u);   // parked in s_next before the resolve
        RAC_PHASE(0);
        raster_tile(sm, s_ctl, s_scan, &s_sc, table, mesh_ids, uniform_mesh, T_CO, K, W, H, maxT, tg, S, stats, large_cnt,
                    large_tri, recs, desc, refs, rgba_out, depth_out, (int)(e / (unsigned)tg.nt), (int)(e % (unsigned)tg.nt),
                    r_next, n_work, s_prefix, buckets, bucket_cap, s_next, need_dstats RAC_PHASE_PASS);
        RAC_PHASE(11);
        if (threadIdx.x == 0) asm volatile("cp.async.wait_all;" ::: "memory");
        __syncthreads();                  // the bin's shared memory is free; thread 0's copy has landed
#ifdef RAC_PHASE_TIMING
        if (threadIdx.x == 0) {           // cycles per load class: u64 slots 14..29 of WorkCtl::pad
            int k = 0;
            for (int step = kBuckets / 2; step > 0; step >>= 1)
                if (s_prefix[k + step] <= r) k += step;
            atomicAdd(reinterpret_cast<unsigned long long*>(ctl->pad) + 14 + k, (unsigned long long)(clock64() - t_bin0));
        }
#endif
        r = s_next[0]; e = s_next[1];     // (s_next is rewritten only after the next bin's coverage barriers)
        RAC_PHASE(12);
    }
#ifdef RAC_PHASE_TIMING
    if (threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicAdd(&g_tl[10], t - g_tl[15]);
    }
#endif
    RAC_TL(2);
}
