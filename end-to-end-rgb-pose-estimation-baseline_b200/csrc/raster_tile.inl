// raster_tile.inl — raster_kernel: one CTA per (image, 32x16-px tile) work item.  Spec: RS-C, RS-Z, RS-R in raster.cuh.
// Included by raster.cu (inside namespace rac).

// Optional phase timeline (build with -DRAC_PHASE_TIMING): thread 0 of every CTA accumulates clock64() deltas per
// phase into WorkCtl::pad (read back from the workspace by profiles/phase_timing.py).  Off in the product build.
#ifdef RAC_PHASE_TIMING
#define RAC_PHASE(idx)                                                                                          \
    do {                                                                                                        \
        if (threadIdx.x == 0) {                                                                                 \
            const long long _t = clock64();                                                                     \
            atomicAdd(reinterpret_cast<unsigned long long*>(g_phase_ctl->pad) + (idx), (unsigned long long)(_t - g_phase_t)); \
            g_phase_t = _t;                                                                                     \
        }                                                                                                       \
    } while (0)
#define RAC_COUNT(idx, v) atomicAdd(reinterpret_cast<unsigned long long*>(g_phase_ctl->pad) + (idx), (unsigned long long)(v))
#define RAC_PHASE_ARGS , WorkCtl* g_phase_ctl, long long& g_phase_t
#define RAC_PHASE_PASS , g_phase_ctl, g_phase_t
#else
#define RAC_PHASE(idx) do { } while (0)
#define RAC_COUNT(idx, v) do { } while (0)
#define RAC_PHASE_ARGS
#define RAC_PHASE_PASS
#endif

constexpr int kRasterThreads = 256;
constexpr int kRasterRegs = 56;       // 4 CTAs/SM leave 8 K registers: one resident gt_norm CTA (256 threads x 32 regs) per SM
constexpr int kQueueCap = 1024;       // triangle queue per flush
constexpr int kSmallPix = 16;         // bbox∩tile pixels handled by one thread; larger -> one warp
constexpr int kTilePix = kTileW * kTileH;

__device__ __forceinline__ void smem_min64(unsigned long long* addr, unsigned long long v) {
    unsigned long long old = *(volatile unsigned long long*)addr;
    while (v < old) {
        const unsigned long long assumed = old;
        old = atomicCAS(addr, assumed, v);
        if (old == assumed) break;
    }
}

// bbox (in pixels) of the triangle clipped to the tile [x_lo, x_hi] x [y_lo, y_hi]
__device__ __forceinline__ void tri_bbox(const TriGeom& g, int x_lo, int x_hi, int y_lo, int y_hi,
                                         int& bx0, int& bx1, int& by0, int& by1) {
    const int xmin = min(g.X0, min(g.X1, g.X2)), xmax = max(g.X0, max(g.X1, g.X2));
    const int ymin = min(g.Y0, min(g.Y1, g.Y2)), ymax = max(g.Y0, max(g.Y1, g.Y2));
    bx0 = max(ceil_div256(xmin - 224), x_lo);
    bx1 = min(floor_div256(xmax - 32), x_hi);
    by0 = max(ceil_div256(ymin - 224), y_lo);
    by1 = min(floor_div256(ymax - 32), y_hi);
}

// RS-C + RS-Z for one triangle over its bbox∩tile, `nl` lanes cooperating (lane in [0,nl)).
// I = int for binned triangles (extent <= 64 px: every product < 2^29), long long for large ones.
// keys: [4][kTilePix] sample-major visibility buffer of this tile (origin tile_x0, tile_y0).
template <typename I>
__device__ __forceinline__ void raster_tri(const TriGeom& g, unsigned tri, int bx0, int bx1, int by0, int by1,
                                           int lane, int nl, unsigned long long* keys, int tile_x0, int tile_y0,
                                           float c0, float c1) {
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
    for (int i = lane; i < n; i += nl) {
        const int ry = i / w;
        const int px = bx0 + (i - ry * w), py = by0 + ry;
        const int ox = px * kSubpix, oy = py * kSubpix;
        const I E0 = dx0 * (I)(oy - g.Y0) - dy0 * (I)(ox - g.X0) + b0;
        const I E1 = dx1 * (I)(oy - g.Y1) - dy1 * (I)(ox - g.X1) + b1;
        const I E2 = dx2 * (I)(oy - g.Y2) - dy2 * (I)(ox - g.X2) + b2;
        const int pidx = (py - tile_y0) * kTileW + (px - tile_x0);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const I e0 = E0 + dx0 * (I)SY[s] - dy0 * (I)SX[s];
            const I e1 = E1 + dx1 * (I)SY[s] - dy1 * (I)SX[s];
            const I e2 = E2 + dx2 * (I)SY[s] - dy2 * (I)SX[s];
            if ((e0 | e1 | e2) >= 0) {
                const float fxr = (float)(ox + SX[s] - g.X0), fyr = (float)(oy + SY[s] - g.Y0);
                const float iz = __fmaf_rn(g.gy, fyr, __fmaf_rn(g.gx, fxr, g.iz0));
                const float d = __fmaf_rn(-c1, iz, c0);
                if (d >= 0.0f && d <= 1.0f) {
                    const unsigned d24 = (unsigned)__float2int_rn(__fmul_rn(d, 16777215.0f));
                    if (d24 < kDepthMax)
                        smem_min64(&keys[s * kTilePix + pidx], ((unsigned long long)d24 << 32) | tri);
                }
            }
        }
    }
}

// Rasterise the `qn` queued triangles into the tile's visibility buffer: triangles whose bbox∩tile is at most
// kSmallPix pixels take one thread each, the others are deferred to one warp each.  Called by all threads.
template <typename I>
__device__ __forceinline__ void process_queue(const int* queue, int qn, int* medq, int* med_count,
                                              const int4* __restrict__ img_recs, unsigned long long* keys,
                                              int x_lo, int x_hi, int y_lo, int y_hi, float c0, float c1) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < qn; k += kRasterThreads) {
        const unsigned tri = (unsigned)queue[k];
        const int4* r = img_recs + 3 * (size_t)tri;
        const TriGeom g = unpack_geom(r[0], r[1], r[2]);
        int bx0, bx1, by0, by1;
        tri_bbox(g, x_lo, x_hi, y_lo, y_hi, bx0, bx1, by0, by1);
        const int npix = (bx1 - bx0 + 1) * (by1 - by0 + 1);
        if (bx0 > bx1 || by0 > by1) continue;
        if (npix <= kSmallPix)
            raster_tri<I>(g, tri, bx0, bx1, by0, by1, 0, 1, keys, x_lo, y_lo, c0, c1);
        else
            medq[atomicAdd(med_count, 1)] = (int)tri;
    }
    __syncthreads();
    const int mn = *med_count;
    for (int k = warp; k < mn; k += kRasterThreads / 32) {
        const unsigned tri = (unsigned)medq[k];
        const int4* r = img_recs + 3 * (size_t)tri;
        const TriGeom g = unpack_geom(r[0], r[1], r[2]);
        int bx0, bx1, by0, by1;
        tri_bbox(g, x_lo, x_hi, y_lo, y_hi, bx0, bx1, by0, by1);
        raster_tri<I>(g, tri, bx0, bx1, by0, by1, lane, 32, keys, x_lo, y_lo, c0, c1);
    }
    __syncthreads();
    if (tid == 0) *med_count = 0;
    __syncthreads();
}

// Binned triangles (int32-exact), flattened so that lanes stay busy: each thread stages one queued triangle
// (edge equations relative to its clipped bbox origin, depth plane, id) in shared memory and lists the bbox
// pixels it must visit; the warp then walks the concatenated (triangle, pixel) items of its 32 triangles, 32 per
// step, so a 1-pixel triangle and a 16-pixel triangle in the same warp cost 17 items, not 2 x 16.
// Triangles with more than kSmallPix bbox pixels take the warp-per-triangle path.
constexpr int kTriStride = 20;   // words per staged triangle (16 used; 80 B keeps LDS.128 aligned and conflict-free)

constexpr size_t kRasterSmemBytes = (size_t)kTilePix * 32 + sizeof(int) * 2 * kQueueCap +
                                    sizeof(int) * kRasterThreads * kTriStride + sizeof(unsigned short) * kRasterThreads * kSmallPix;

__device__ __forceinline__ void process_queue_flat(const int* queue, int qn, int* medq, int* med_count,
                                                   const int4* __restrict__ img_recs, unsigned long long* keys,
                                                   int* s_tri, unsigned short* s_items,
                                                   int x_lo, int x_hi, int y_lo, int y_hi, float c0, float c1
                                                   RAC_PHASE_ARGS) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int SX[4] = {96, 224, 32, 160};
    constexpr int SY[4] = {32, 96, 160, 224};
    unsigned short* items = s_items + warp * (32 * kSmallPix);
    for (int qb = 0; qb < qn; qb += kRasterThreads) {
        const int k = qb + tid;
        int n = 0, w = 0, h = 0;
        if (k < qn) {
            const unsigned tri = (unsigned)queue[k];
            const int4* r = img_recs + 3 * (size_t)tri;
            const int4 q0 = r[0], q1 = r[1], q2 = r[2];
            const TriGeom g = unpack_geom(q0, q1, q2);
            int bx0, bx1, by0, by1;
            tri_bbox(g, x_lo, x_hi, y_lo, y_hi, bx0, bx1, by0, by1);
            if (bx0 <= bx1 && by0 <= by1) {
                w = bx1 - bx0 + 1; h = by1 - by0 + 1;
                if (w * h <= kSmallPix) {
                    n = w * h;
                    const int dx0 = g.X1 - g.X0, dy0 = g.Y1 - g.Y0;
                    const int dx1 = g.X2 - g.X1, dy1 = g.Y2 - g.Y1;
                    const int dx2 = g.X0 - g.X2, dy2 = g.Y0 - g.Y2;
                    const int b0 = (dy0 < 0 || (dy0 == 0 && dx0 > 0)) ? 0 : -1;
                    const int b1 = (dy1 < 0 || (dy1 == 0 && dx1 > 0)) ? 0 : -1;
                    const int b2 = (dy2 < 0 || (dy2 == 0 && dx2 > 0)) ? 0 : -1;
                    const int ox = bx0 * kSubpix, oy = by0 * kSubpix;
                    int4* slot = reinterpret_cast<int4*>(s_tri + tid * kTriStride);
                    const int C0 = dx0 * (oy - g.Y0) - dy0 * (ox - g.X0) + b0;
                    const int C1 = dx1 * (oy - g.Y1) - dy1 * (ox - g.X1) + b1;
                    const int C2 = dx2 * (oy - g.Y2) - dy2 * (ox - g.X2) + b2;
                    slot[0] = make_int4(-dy0, dx0, C0, -dy1);
                    slot[1] = make_int4(dx1, C1, -dy2, dx2);
                    slot[2] = make_int4(C2, ox - g.X0, oy - g.Y0, __float_as_int(g.gx));
                    slot[3] = make_int4(__float_as_int(g.gy), __float_as_int(g.iz0), (int)tri,
                                        (by0 - y_lo) * kTileW + (bx0 - x_lo));
                } else {
                    medq[atomicAdd(med_count, 1)] = (int)tri;
                }
            }
        }
        int incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        if (n) {
            int idx = incl - n;
            for (int dy = 0; dy < h; ++dy)
                for (int dx = 0; dx < w; ++dx) items[idx++] = (unsigned short)(lane | (dx << 5) | (dy << 9));
        }
        __syncwarp();
        for (int j = lane; j < total; j += 32) {
            const int code = items[j];
            const int dxp = (code >> 5) & 15, dyp = code >> 9;
            const int4* slot = reinterpret_cast<const int4*>(s_tri + ((warp << 5) + (code & 31)) * kTriStride);
            const int4 a = slot[0], b = slot[1], c = slot[2], d = slot[3];
            const int sx = dxp * kSubpix, sy = dyp * kSubpix;
            const float gx = __int_as_float(c.w), gy = __int_as_float(d.x), iz0 = __int_as_float(d.y);
            const int pidx = d.w + dyp * kTileW + dxp;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int ux = sx + SX[s], uy = sy + SY[s];
                const int e0 = a.z + a.x * ux + a.y * uy;
                const int e1 = b.y + a.w * ux + b.x * uy;
                const int e2 = c.x + b.z * ux + b.w * uy;
                if ((e0 | e1 | e2) >= 0) {
                    const float fxr = (float)(c.y + ux), fyr = (float)(c.z + uy);
                    const float iz = __fmaf_rn(gy, fyr, __fmaf_rn(gx, fxr, iz0));
                    const float dd = __fmaf_rn(-c1, iz, c0);
                    if (dd >= 0.0f && dd <= 1.0f) {
                        const unsigned d24 = (unsigned)__float2int_rn(__fmul_rn(dd, 16777215.0f));
                        if (d24 < kDepthMax)
                            smem_min64(&keys[s * kTilePix + pidx], ((unsigned long long)d24 << 32) | (unsigned)d.z);
                    }
                }
            }
        }
        __syncwarp();
    }
    RAC_PHASE(2);
    __syncthreads();
    RAC_PHASE(3);
    const int mn = *med_count;
    for (int k = warp; k < mn; k += kRasterThreads / 32) {
        const unsigned tri = (unsigned)medq[k];
        const int4* r = img_recs + 3 * (size_t)tri;
        const TriGeom g = unpack_geom(r[0], r[1], r[2]);
        int bx0, bx1, by0, by1;
        tri_bbox(g, x_lo, x_hi, y_lo, y_hi, bx0, bx1, by0, by1);
        raster_tri<int>(g, tri, bx0, bx1, by0, by1, lane, 32, keys, x_lo, y_lo, c0, c1);
    }
    __syncthreads();
    RAC_PHASE(4);
    if (tid == 0) *med_count = 0;
    __syncthreads();
}

struct TileAcc {
    unsigned qmax, dcount;
    float dmax;
    double dsum, dsumsq;
};

// Everything one CTA does for tile `tile` of image `b`: gather -> coverage/depth -> shade/resolve -> RGBA8 + depth
// stores -> per-image accumulators.  smem_raw: keys [4][kTilePix] u64, queue [kQueueCap], medq [kQueueCap],
// staged triangles [kRasterThreads][kTriStride], item lists [8][32*kSmallPix] u16.
__device__ __forceinline__ void raster_tile(unsigned char* smem_raw, int* s_ctl, unsigned* s_scan,
                                            const MeshDev* __restrict__ table, const int* __restrict__ mesh_ids,
                                            int uniform_mesh, const float* __restrict__ T_CO, const float* __restrict__ K,
                                            int W, int H, int maxT, const TileGrid& tg, const SceneConsts& S,
                                            ImgStats* __restrict__ stats, const unsigned* __restrict__ large_tri,
                                            const int4* __restrict__ recs, const unsigned* __restrict__ desc,
                                            const unsigned char* __restrict__ refs, unsigned* __restrict__ rgba_out,
                                            float* __restrict__ depth_out, int b, int tile RAC_PHASE_ARGS) {
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw);           // [4][kTilePix]
    int* queue = reinterpret_cast<int*>(keys + 4 * kTilePix);                              // [kQueueCap]
    int* medq = queue + kQueueCap;                                                         // [kQueueCap]
    int* s_tri = medq + kQueueCap;                                                         // [kRasterThreads][kTriStride]
    unsigned short* s_items = reinterpret_cast<unsigned short*>(s_tri + kRasterThreads * kTriStride);  // [8][32*kSmallPix]
    int* q_count = s_ctl, *med_count = s_ctl + 1;
    unsigned* s_dcnt = s_scan;                      // [kRasterThreads] descriptor counts
    unsigned* s_dscan = s_scan + kRasterThreads;    // [kRasterThreads] exclusive scan
    unsigned* s_doff = s_scan + 2 * kRasterThreads; // [kRasterThreads] offset into the ref region
    unsigned* s_wsum = s_scan + 3 * kRasterThreads; // [8]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx = tile % tg.ntx, ty = tile / tg.ntx;
    const int x_lo = tx * kTileW, y_lo = ty * kTileH;
    const int x_hi = min(x_lo + kTileW, W) - 1, y_hi = min(y_lo + kTileH, H) - 1;
    const MeshDev& mesh = table[mesh_ids ? mesh_ids[b] : uniform_mesh];
    const int4* img_recs = recs + 3 * (size_t)b * maxT;
    const unsigned n_large = stats[b].nlarge;   // written by setup_kernel (previous launch)
    const int ncta = tg.binned ? (mesh.T + kSetupThreads - 1) / kSetupThreads : 0;

    if (tid == 0) { *q_count = 0; *med_count = 0; }
    bool inited = false;   // CTA-uniform
    __syncthreads();

    // ---- binned triangles: descriptors of this tile, kRasterThreads at a time ----
    const unsigned* dtile = desc + ((size_t)b * tg.nt + tile) * tg.ncta_max;
    for (int dbase = 0; dbase < ncta; dbase += kRasterThreads) {
        const int j = dbase + tid;
        const unsigned d = j < ncta ? dtile[j] : 0u;
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
        for (int w = 0; w < kRasterThreads / 32; ++w) {
            const unsigned v = s_wsum[w];
            if (w < warp) wbase += v;
            total += v;
        }
        s_dcnt[tid] = c; s_dscan[tid] = wbase + incl - c; s_doff[tid] = d >> 16;
        __syncthreads();
        const int nd = min(kRasterThreads, ncta - dbase);
        for (unsigned base = 0; base < total; base += kQueueCap) {
            // expand refs [base, base + kQueueCap) of this descriptor group into the queue, one warp per descriptor
            for (int jj = warp; jj < nd; jj += kRasterThreads / 32) {
                const unsigned cj = s_dcnt[jj];
                if (cj == 0) continue;
                const unsigned sj = s_dscan[jj];
                const unsigned lo = max(sj, base), hi = min(sj + cj, base + (unsigned)kQueueCap);
                const unsigned char* region = refs + (((size_t)b * tg.ncta_max + dbase + jj) << 10) + s_doff[jj];
                for (unsigned k = lo + lane; k < hi; k += 32)
                    queue[k - base] = (dbase + jj) * kSetupThreads + (int)region[k - sj];
            }
            if (!inited) {
                for (int k = tid; k < 4 * kTilePix; k += kRasterThreads) keys[k] = kKeyEmpty;
                inited = true;
            }
            __syncthreads();
            RAC_PHASE(1);
            const int qn = (int)min((unsigned)kQueueCap, total - base);
            process_queue_flat(queue, qn, medq, med_count, img_recs, keys, s_tri, s_items, x_lo, x_hi, y_lo, y_hi, S.c0, S.c1 RAC_PHASE_PASS);
        }
    }

    // ---- large triangles: every tile scans the image's large list ----
    for (unsigned base = 0; base < n_large; base += kRasterThreads) {
        const unsigned i = base + tid;
        if (i < n_large) {
            const unsigned tri = large_tri[(size_t)b * maxT + i];
            const int4* r = img_recs + 3 * (size_t)tri;
            const int4 q0 = r[0], q1 = r[1];
            const int xmin = min(q0.x, min(q0.z, q1.x)), xmax = max(q0.x, max(q0.z, q1.x));
            const int ymin = min(q0.y, min(q0.w, q1.y)), ymax = max(q0.y, max(q0.w, q1.y));
            if (ceil_div256(xmin - 224) <= x_hi && floor_div256(xmax - 32) >= x_lo &&
                ceil_div256(ymin - 224) <= y_hi && floor_div256(ymax - 32) >= y_lo)
                queue[atomicAdd(q_count, 1)] = (int)tri;
        }
        __syncthreads();
        const int qn = *q_count;
        const bool last = base + kRasterThreads >= n_large;
        if (qn > 0 && (last || qn > kQueueCap - kRasterThreads)) {
            if (!inited) {
                for (int k = tid; k < 4 * kTilePix; k += kRasterThreads) keys[k] = kKeyEmpty;
                inited = true;
                __syncthreads();
            }
            process_queue<long long>(queue, qn, medq, med_count, img_recs, keys, x_lo, x_hi, y_lo, y_hi, S.c0, S.c1);
            if (tid == 0) *q_count = 0;
            __syncthreads();
        }
    }

    RAC_PHASE(5);
    // ---- resolve + shade + write, and per-image accumulators ----
    const int tw = x_hi - x_lo + 1, th = y_hi - y_lo + 1;
    TileAcc acc;
    acc.qmax = 0; acc.dcount = 0; acc.dmax = 0.0f; acc.dsum = 0.0; acc.dsumsq = 0.0;
    const size_t img_pix = (size_t)b * W * H;
    if (!inited) {
        for (int p = tid; p < kTilePix; p += kRasterThreads) {
            const int lx = p & (kTileW - 1), ly = p / kTileW;
            if (lx < tw && ly < th) {
                const size_t o = img_pix + (size_t)(y_lo + ly) * W + (x_lo + lx);
                rgba_out[o] = 0u;
                depth_out[o] = 0.0f;
            }
        }
    } else {
        // Flattened shading.  R1: every pixel finds the distinct triangles among its 4 samples and lists one
        // (pixel, sample) item per distinct triangle; R2: the items are shaded 256 at a time (no lane idles on
        // background or waits for a neighbour's 4-triangle edge pixel) and the RGBA8 colour replaces the triangle
        // id in the key's low word; R3: every pixel sums its samples' colours through the representatives.
        const CamPose cam = load_campose(T_CO, K, b);
        const ShadeCtx sc = make_shade_ctx(mesh, cam, S);
        unsigned short* sitems = reinterpret_cast<unsigned short*>(queue);   // [4 * kTilePix] aliases queue + medq
        unsigned* key_lo = reinterpret_cast<unsigned*>(keys);
        constexpr int kPer = kTilePix / kRasterThreads;
        unsigned info[kPer];     // per pixel: rep[s] (2 bits each) | covered mask << 8
        unsigned d24s[kPer];     // depth of sample 0
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            const int p = tid + i * kRasterThreads;
            const int lx = p & (kTileW - 1), ly = p / kTileW;
            unsigned long long k[4];
#pragma unroll
            for (int sidx = 0; sidx < 4; ++sidx)
                k[sidx] = (lx < tw && ly < th) ? keys[sidx * kTilePix + p] : kKeyEmpty;
            unsigned reps = 0, mask = 0, fresh = 0;
#pragma unroll
            for (int sidx = 0; sidx < 4; ++sidx) {
                if (k[sidx] != kKeyEmpty) {
                    mask |= 1u << sidx;
                    unsigned rep = sidx;
#pragma unroll
                    for (int j = sidx - 1; j >= 0; --j)
                        if (k[j] != kKeyEmpty && (unsigned)k[j] == (unsigned)k[sidx]) rep = j;
                    reps |= rep << (2 * sidx);
                    if (rep == (unsigned)sidx) fresh |= 1u << sidx;
                }
            }
            info[i] = reps | (mask << 8);
            d24s[i] = (mask & 1u) ? (unsigned)(k[0] >> 32) : kDepthMax;
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
            for (int sidx = 0; sidx < 4; ++sidx)
                if (fresh & (1u << sidx)) sitems[base++] = (unsigned short)(p | (sidx << 10));
        }
        RAC_PHASE(6);
        __syncthreads();
        RAC_PHASE(7);
        const int nitems = *q_count;
        for (int it = tid; it < nitems; it += kRasterThreads) {
            const int code = sitems[it];
            const int p = code & 1023, sidx = code >> 10;
            const unsigned tri = key_lo[2 * (sidx * kTilePix + p)];
            key_lo[2 * (sidx * kTilePix + p)] = shade(sc, img_recs + 3 * (size_t)tri, tri, x_lo + (p & (kTileW - 1)), y_lo + p / kTileW);
        }
        RAC_PHASE(8);
        __syncthreads();
        RAC_PHASE(9);
        if (tid == 0) *q_count = 0;
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            const int p = tid + i * kRasterThreads;
            const int lx = p & (kTileW - 1), ly = p / kTileW;
            if (lx >= tw || ly >= th) continue;
            const unsigned mask = info[i] >> 8;
            unsigned sum_r = 0, sum_g = 0, sum_b = 0, sum_a = 0;
#pragma unroll
            for (int sidx = 0; sidx < 4; ++sidx) {
                if (mask & (1u << sidx)) {
                    const unsigned rep = (info[i] >> (2 * sidx)) & 3u;
                    const unsigned col = key_lo[2 * (rep * kTilePix + p)];
                    sum_r += col & 0xFFu; sum_g += (col >> 8) & 0xFFu;
                    sum_b += (col >> 16) & 0xFFu; sum_a += col >> 24;
                }
            }
            const unsigned r8 = (sum_r + 2u) >> 2, g8 = (sum_g + 2u) >> 2, b8 = (sum_b + 2u) >> 2, a8 = (sum_a + 2u) >> 2;
            const size_t o = img_pix + (size_t)(y_lo + ly) * W + (x_lo + lx);
            rgba_out[o] = r8 | (g8 << 8) | (b8 << 16) | (a8 << 24);
            float depth = 0.0f;
            if (mask & 1u) depth = depth_readback(d24s[i], S);
            depth_out[o] = depth;
            acc.qmax = max(acc.qmax, max(r8, max(g8, b8)));
            acc.dmax = fmaxf(acc.dmax, depth);
            if (depth > 0.01f) { acc.dcount++; acc.dsum += (double)depth; acc.dsumsq += (double)depth * (double)depth; }
        }
    }
    RAC_PHASE(10);
    // block reduction -> one atomic per accumulator per CTA
    __shared__ double red_d[2][kRasterThreads / 32];
    __shared__ unsigned red_u[3][kRasterThreads / 32];
    acc.qmax = warp_max(acc.qmax);
    acc.dcount = (unsigned)__reduce_add_sync(0xffffffffu, acc.dcount);
    const unsigned dmax_bits = warp_max(__float_as_uint(acc.dmax));
    acc.dsum = warp_sum(acc.dsum); acc.dsumsq = warp_sum(acc.dsumsq);
    if (lane == 0) {
        red_u[0][warp] = acc.qmax; red_u[1][warp] = acc.dcount; red_u[2][warp] = dmax_bits;
        red_d[0][warp] = acc.dsum; red_d[1][warp] = acc.dsumsq;
    }
    __syncthreads();
    if (tid == 0) {
        unsigned qm = 0, dc = 0, dm = 0;
        double a = 0, c = 0;
        for (int w = 0; w < kRasterThreads / 32; ++w) {
            qm = max(qm, red_u[0][w]); dc += red_u[1][w]; dm = max(dm, red_u[2][w]);
            a += red_d[0][w]; c += red_d[1][w];
        }
        ImgStats* st = stats + b;
        if (qm) atomicMax(&st->qmax, qm);
        if (dc) { atomicAdd(&st->dcount, dc); atomicAdd(&st->dsum, a); atomicAdd(&st->dsumsq, c); }
        if (dm) atomicMax(&st->dmax_bits, dm);
    }
}

// Persistent CTAs (SMs x resident CTAs) pop non-empty tiles from the work list setup_kernel built, image-major
// through img_base (item r belongs to image b with img_base[b] <= r < img_base[b+1]).  The claim for the NEXT item
// is issued before the current tile is rasterised, so the atomic's round trip to L2 is hidden behind the tile.
__global__ void __maxnreg__(kRasterRegs)
raster_kernel(const MeshDev* __restrict__ table, const int* __restrict__ mesh_ids, int uniform_mesh,
              const float* __restrict__ T_CO, const float* __restrict__ K, int W, int H, int maxT, TileGrid tg,
              SceneConsts S, ImgStats* __restrict__ stats, const unsigned* __restrict__ large_tri,
              const int4* __restrict__ recs, const unsigned* __restrict__ desc, const unsigned char* __restrict__ refs,
              WorkCtl* __restrict__ ctl, const unsigned* __restrict__ worklist, const unsigned* __restrict__ img_base,
              unsigned* __restrict__ rgba_out, float* __restrict__ depth_out, int B) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_ctl[2];
    __shared__ unsigned s_item;
    __shared__ unsigned s_scan[3 * kRasterThreads + kRasterThreads / 32];
    const unsigned n_work = img_base[B];    // final: written by setup_kernel (previous launch)
#ifdef RAC_PHASE_TIMING
    WorkCtl* g_phase_ctl = ctl;
    long long g_phase_t = clock64();
#endif
    RAC_TL(2);
#ifdef RAC_PHASE_TIMING
    if (threadIdx.x == 0) {      // late starters: CTAs that became resident more than 10 us after the first one
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t > g_tl[4] + 10000ull) atomicAdd(&g_tl[8], 1ull);
    }
#endif
    if (threadIdx.x == 0) s_item = atomicAdd(&ctl->next, 1u);
    __syncthreads();
    for (;;) {
        const unsigned r = s_item;
        if (r >= n_work) break;
        unsigned r_next = 0;
        if (threadIdx.x == 0) r_next = atomicAdd(&ctl->next, 1u);   // consumed after the tile
        int lo = 0, hi = B;          // img_base[lo] <= r < img_base[hi]
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (img_base[mid] <= r) lo = mid; else hi = mid;
        }
        const unsigned item = (unsigned)lo * (unsigned)tg.nt + (r - img_base[lo]);
        RAC_PHASE(0);
        raster_tile(smem_raw, s_ctl, s_scan, table, mesh_ids, uniform_mesh, T_CO, K, W, H, maxT, tg, S, stats, large_tri,
                    recs, desc, refs, rgba_out, depth_out, lo, (int)worklist[item] RAC_PHASE_PASS);
        RAC_PHASE(11);
        __syncthreads();                  // every thread has read s_item (and left the tile's shared memory)
        if (threadIdx.x == 0) s_item = r_next;
        __syncthreads();
        RAC_PHASE(12);
    }
    RAC_TL(2);
}
