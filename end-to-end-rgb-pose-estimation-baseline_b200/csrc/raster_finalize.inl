// raster_finalize.inl — streaming normalise + concat kernels (RS-N, data_loaders.py:69-82,158-177).
// Included by raster.cu (inside namespace rac).

// ---------------------------------------------------------------------------------------------
// finalize: normalise + concat / HWC expand
// ---------------------------------------------------------------------------------------------
struct NormConsts {
    float inv_qmax_valid;  // 1 if qmax > 0
    float qmax;
    bool depth_norm;       // apply (d-mean)/std on d > 0.01
    float dmean, dstd;
};

__device__ __forceinline__ NormConsts load_norm(const ImgStats& st, bool want_depth_norm) {
    NormConsts n;
    n.qmax = (float)st.qmax;
    n.inv_qmax_valid = st.qmax > 0 ? 1.0f : 0.0f;
    n.depth_norm = false; n.dmean = 0.f; n.dstd = 1.f;
    if (want_depth_norm) {
        // ref: data_loaders.py:69-77 — two early returns hand back the depth unchanged (Q10)
        const float dmax = __uint_as_float(st.dmax_bits);
        if (dmax > 0.1f && st.dcount > 0) {
            const double mean = st.dsum / (double)st.dcount;
            double var = st.dsumsq / (double)st.dcount - mean * mean;
            if (var < 0.0) var = 0.0;
            const float stdf = (float)sqrt(var);
            if (!(fabsf(stdf) <= 1e-8f)) { n.depth_norm = true; n.dmean = (float)mean; n.dstd = stdf; }
        }
    }
    return n;
}

__device__ __forceinline__ float norm_depth1(float d, const NormConsts& n) {
    if (!n.depth_norm) return d;
    return d > 0.01f ? __fdiv_rn(__fsub_rn(d, n.dmean), n.dstd) : 0.0f;
}
__device__ __forceinline__ float norm_col1(unsigned q, const NormConsts& n) {
    return n.inv_qmax_valid != 0.0f ? __fdiv_rn((float)q, n.qmax) : 0.0f;
}

// Everything the finalize kernels need.  Outputs are optional (nullptr = not requested) except where noted.
struct FinArgs {
    const ImgStats* stats;
    const unsigned* rgba;      // resolved RGBA8 framebuffer (workspace; live tiles only)
    const float* init_hwc;     // rac_assemble_input only: an existing f32 HWC render
    const float* depth;        // metric depth (workspace, live tiles only; or the caller's depth map for rac_assemble_input)
    const float* gt_hwc;       // the gt crop (handled by gt_norm_kernel, kept here for the host plumbing)
    int npix, W, use_norm_depth;
    float* x_nchw;             // (B,C,H,W) concat: [norm depth?, render R, G, B, gt R, G, B]
    float* rgb_hwc_out;        // (B,H,W,3) max-normalised render (what render_batch returns)
    unsigned* rgba_copy;       // (B,H,W) RGBA8 framebuffer copy
    float* depth_out;          // (B,H,W) metric depth for the caller
    int mode, vec;
};

// ImgStats through L2 (ld.global.cg): they were produced by other SMs' atomics.
__device__ __forceinline__ ImgStats load_stats_cg(const ImgStats* p) {
    ImgStats st;
    const int4* src = reinterpret_cast<const int4*>(p);
    int4* dst = reinterpret_cast<int4*>(&st);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = __ldcg(src + i);
    return st;
}

// ---------------------------------------------------------------------------------------------
// live tiles: RGBA8 + depth (workspace, L2-resident) -> normalised outputs
// ---------------------------------------------------------------------------------------------
// The background of every output was zero-filled on the forked stream while the image was being rasterised (the gt
// kernel does it for x, memsets for the others), so after raster_kernel only the non-empty tiles of the work list are
// touched: ~1/4 of the pixels at the bench workload.  One CTA of 128 threads per tile, 4 pixels of one row per
// thread: every row segment of a 32-px tile is one 128-B line per plane.
constexpr int kFinThreads = 128;

__global__ void __launch_bounds__(kFinThreads)
finalize_live_kernel(FinArgs A, TileGrid tg, int H, const unsigned* __restrict__ buckets, size_t bucket_cap, const WorkCtl* __restrict__ ctl) {
    RAC_TL(3);
    __shared__ unsigned s_prefix[kBuckets + 1];
    load_bucket_prefix(ctl, s_prefix);
    __syncthreads();
    const unsigned n_work = s_prefix[kBuckets];
    const int W = A.W, npix = A.npix;
    const int C = A.use_norm_depth ? 7 : 6;
    const int ly = threadIdx.x >> 3, lx = (threadIdx.x & 7) << 2;
    __shared__ float s_lut[256];      // q / qmax for the 256 possible 8-bit values of this image (R9): 256 divisions per
                                      // bin instead of 1536
    for (unsigned r = blockIdx.x; r < n_work; r += gridDim.x) {
        const unsigned e = bucket_entry(s_prefix, buckets, bucket_cap, r);
        const int b = (int)(e / (unsigned)tg.nt), tile = (int)(e % (unsigned)tg.nt);
        const int x = (tile % tg.ntx) * kTileW + lx, y = (tile / tg.ntx) * kTileH + ly;
        // the bin's pixels are requested first: their L2 round trip overlaps the per-image statistics load and the
        // construction of the lookup table (two barriers) instead of following them
        const bool inside = y < H && x < W;
        const size_t pi = (size_t)b * npix + (size_t)y * W + x;
        const int nv = min(4, W - x);
        unsigned c[4] = {0u, 0u, 0u, 0u};
        float d[4] = {0.f, 0.f, 0.f, 0.f};
        const bool want_d = A.depth_out || (A.x_nchw && A.use_norm_depth);
        if (inside) {
            if (A.vec) {
                const uint4 cv = __ldcg(reinterpret_cast<const uint4*>(A.rgba + pi));
                c[0] = cv.x; c[1] = cv.y; c[2] = cv.z; c[3] = cv.w;
                if (want_d) {
                    const float4 dv = __ldcg(reinterpret_cast<const float4*>(A.depth + pi));
                    d[0] = dv.x; d[1] = dv.y; d[2] = dv.z; d[3] = dv.w;
                }
            } else {
                for (int i = 0; i < nv; ++i) { c[i] = __ldcg(A.rgba + pi + i); if (want_d) d[i] = __ldcg(A.depth + pi + i); }
            }
        }
        const ImgStats st = load_stats_cg(A.stats + b);
        const NormConsts n = load_norm(st, A.use_norm_depth != 0);
        __syncthreads();              // previous bin's lookups are done
        for (int q = threadIdx.x; q < 256; q += kFinThreads) s_lut[q] = norm_col1((unsigned)q, n);
        __syncthreads();
        if (!inside) continue;
        float rr[4], gg[4], bb[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            rr[i] = s_lut[c[i] & 0xFFu]; gg[i] = s_lut[(c[i] >> 8) & 0xFFu]; bb[i] = s_lut[(c[i] >> 16) & 0xFFu];
        }
        if (A.vec) {
            if (A.x_nchw) {
                float* xb = A.x_nchw + (size_t)b * C * npix + (size_t)y * W + x;
                int ch = 0;
                if (A.use_norm_depth) {
                    __stcs(reinterpret_cast<float4*>(xb), make_float4(norm_depth1(d[0], n), norm_depth1(d[1], n), norm_depth1(d[2], n), norm_depth1(d[3], n)));
                    ch = 1;
                }
                __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 0) * npix), make_float4(rr[0], rr[1], rr[2], rr[3]));
                __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 1) * npix), make_float4(gg[0], gg[1], gg[2], gg[3]));
                __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 2) * npix), make_float4(bb[0], bb[1], bb[2], bb[3]));
            }
            if (A.rgb_hwc_out) {
                float4* o = reinterpret_cast<float4*>(A.rgb_hwc_out + 3 * pi);
                __stcs(o + 0, make_float4(rr[0], gg[0], bb[0], rr[1]));
                __stcs(o + 1, make_float4(gg[1], bb[1], rr[2], gg[2]));
                __stcs(o + 2, make_float4(bb[2], rr[3], gg[3], bb[3]));
            }
            if (A.rgba_copy) *reinterpret_cast<uint4*>(A.rgba_copy + pi) = make_uint4(c[0], c[1], c[2], c[3]);
            if (A.depth_out) *reinterpret_cast<float4*>(A.depth_out + pi) = make_float4(d[0], d[1], d[2], d[3]);
        } else {
            for (int i = 0; i < nv; ++i) {
                if (A.x_nchw) {
                    float* xb = A.x_nchw + (size_t)b * C * npix + (size_t)y * W + x + i;
                    int ch = 0;
                    if (A.use_norm_depth) { xb[0] = norm_depth1(d[i], n); ch = 1; }
                    xb[(size_t)(ch + 0) * npix] = rr[i]; xb[(size_t)(ch + 1) * npix] = gg[i]; xb[(size_t)(ch + 2) * npix] = bb[i];
                }
                if (A.rgb_hwc_out) { float* o = A.rgb_hwc_out + 3 * (pi + i); o[0] = rr[i]; o[1] = gg[i]; o[2] = bb[i]; }
                if (A.rgba_copy) A.rgba_copy[pi + i] = c[i];
                if (A.depth_out) A.depth_out[pi + i] = d[i];
            }
        }
    }
    RAC_TL(3);
}

// ---------------------------------------------------------------------------------------------
// rac_assemble_input: prepare_model_input on an existing f32 render (data_loaders.py:158-177, render + depth planes)
// ---------------------------------------------------------------------------------------------
constexpr int kFinChunkPix = 2048;

// grid (chunks, B); one thread = 4 consecutive pixels (vec) or 1 pixel
__global__ void __launch_bounds__(256)
assemble_render_kernel(FinArgs A) {
    const int b = blockIdx.y;
    const ImgStats st = load_stats_cg(A.stats + b);
    const NormConsts n = load_norm(st, A.use_norm_depth != 0);
    const int npix = A.npix;
    const int C = A.use_norm_depth ? 7 : 6;
    const int p0 = blockIdx.x * kFinChunkPix, p1 = min(npix, p0 + kFinChunkPix);
    float* xb = A.x_nchw + (size_t)b * C * npix;
    const float* init = A.init_hwc + 3 * (size_t)b * npix;
    const float* dep = A.depth ? A.depth + (size_t)b * npix : nullptr;
    const int ch = A.use_norm_depth ? 1 : 0;
    if (A.vec) {
        for (int q = (p0 >> 2) + (int)threadIdx.x; 4 * q < p1; q += (int)blockDim.x) {
            if (A.use_norm_depth) {
                const float4 d = __ldcs(reinterpret_cast<const float4*>(dep) + q);
                __stcs(reinterpret_cast<float4*>(xb) + q,
                       make_float4(norm_depth1(d.x, n), norm_depth1(d.y, n), norm_depth1(d.z, n), norm_depth1(d.w, n)));
            }
            const float4* s = reinterpret_cast<const float4*>(init) + 3 * (size_t)q;
            const float4 a = __ldcs(s), c = __ldcs(s + 1), e = __ldcs(s + 2);
            __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 0) * npix) + q, make_float4(a.x, a.w, c.z, e.y));
            __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 1) * npix) + q, make_float4(a.y, c.x, c.w, e.z));
            __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 2) * npix) + q, make_float4(a.z, c.y, e.x, e.w));
        }
    } else {
        for (int p = p0 + (int)threadIdx.x; p < p1; p += (int)blockDim.x) {
            if (A.use_norm_depth) xb[p] = norm_depth1(dep[p], n);
            xb[(size_t)(ch + 0) * npix + p] = init[3 * (size_t)p];
            xb[(size_t)(ch + 1) * npix + p] = init[3 * (size_t)p + 1];
            xb[(size_t)(ch + 2) * npix + p] = init[3 * (size_t)p + 2];
        }
    }
}

// normalize_depth statistics of an existing depth map (rac_assemble_input with use_norm_depth; the render paths get
// them from raster_kernel's tile epilogue)
__global__ void __launch_bounds__(256)
depth_stats_kernel(ImgStats* __restrict__ stats, const float* __restrict__ depth, int npix, int chunks) {
    const int b = blockIdx.y;
    const int per = (npix + chunks - 1) / chunks;
    const int p0 = blockIdx.x * per, p1 = min(p0 + per, npix);
    double ds = 0, dss = 0;
    unsigned dc = 0;
    float dmax = 0.f;
    const float* d = depth + (size_t)b * npix;
    for (int i = p0 + threadIdx.x; i < p1; i += blockDim.x) {
        const float v = d[i];
        dmax = fmaxf(dmax, v);
        if (v > 0.01f) { dc++; ds += (double)v; dss += (double)v * v; }
    }
    __shared__ double rd[2][8];
    __shared__ unsigned ru[2][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    ds = warp_sum(ds); dss = warp_sum(dss);
    dc = (unsigned)__reduce_add_sync(0xffffffffu, dc);
    const unsigned dm = warp_max(__float_as_uint(fmaxf(dmax, 0.0f)));
    if (lane == 0) { rd[0][warp] = ds; rd[1][warp] = dss; ru[0][warp] = dc; ru[1][warp] = dm; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double e = 0, f = 0; unsigned u = 0, m = 0;
        for (int w = 0; w < (int)blockDim.x / 32; ++w) { e += rd[0][w]; f += rd[1][w]; u += ru[0][w]; m = max(m, ru[1][w]); }
        ImgStats* st = stats + b;
        if (u) { atomicAdd(&st->dcount, u); atomicAdd(&st->dsum, e); atomicAdd(&st->dsumsq, f); }
        if (m) atomicMax(&st->dmax_bits, m);
    }
}

// ---------------------------------------------------------------------------------------------
// gt crop: normalize_img (data_loaders.py:79-82) + HWC -> CHW into channels [ch0, ch0+3) of x (data_loaders.py:163,174)
// ---------------------------------------------------------------------------------------------
// One thread-block cluster of kGtCluster CTAs per image.  Every CTA first zero-fills its slice of the `nzero` leading
// planes of x (the render / depth background; finalize_live_kernel later overwrites the non-empty tiles), then streams
// its slice of the crop once from HBM (sum / sum of squares: f32 over the 12 values of a pixel quad, f64 across quads),
// the cluster combines the partials through distributed shared memory in rank order (deterministic, no atomics), and
// every CTA streams its slice again — an L2 hit, it read the same lines a few microseconds ago — and writes the three
// normalised planes with 128-bit streaming stores.  mean and 1/std are formed in f64; the per-value affine map runs
// as one f32 fma (absolute error <= 2e-7 on values of magnitude ~1, the size of the f32 rounding of the input itself).
// HBM traffic = 12 B/px read + 12 B/px (+ 4 B/px per zeroed plane) written: the algorithmic minimum of the concat.
// The kernel does not depend on the render, so the render paths launch it on a forked stream where it overlaps
// setup_kernel / raster_kernel (which are issue-bound and leave HBM idle).
constexpr int kGtCluster = 8;
constexpr int kGtThreads = 256;

__global__ void __launch_bounds__(kGtThreads, 8)
gt_norm_kernel(const float* __restrict__ gt_hwc, float* __restrict__ x_nchw, int npix, int C, int ch0, int nzero, int vec, int B,
               const unsigned* __restrict__ live_loads, int W, int ntx, int nt) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ double s_warp[2][kGtThreads / 32];
    __shared__ double s_part[2];      // this CTA's partial sums (read by the whole cluster through DSMEM)
    __shared__ float s_aff[2];
    const int rank = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    RAC_TL(0);
    for (int b = blockIdx.y; b < B; b += gridDim.y) {     // cluster-uniform
    const float* g = gt_hwc + 3 * (size_t)b * npix;
    float* x0 = x_nchw + (size_t)b * C * (size_t)npix;
    float* xb = x0 + (size_t)ch0 * (size_t)npix;
    // slice of this CTA, in quads of 4 pixels (vec) or in pixels
    const int nunit = vec ? npix >> 2 : npix;
    const int u0 = (int)((long long)nunit * rank / kGtCluster), u1 = (int)((long long)nunit * (rank + 1) / kGtCluster);
    if (vec && live_loads && nzero > 0) {
        // live_loads (setup_kernel's per-bin load estimates, final before this kernel starts; W % 4 == 0): the quads
        // of non-empty bins are left alone — finalize_live_kernel writes every pixel of those bins, so the two kernels
        // touch disjoint parts of the render planes and need no ordering between them
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        const unsigned* ll = live_loads + (size_t)b * nt;
        for (int q = u0 + tid; q < u1; q += kGtThreads) {
            const int p = q << 2, y = p / W, x = p - y * W;
            if (__ldg(ll + (y / kTileH) * ntx + x / kTileW) != 0u) continue;
            for (int pl = 0; pl < nzero; ++pl) __stcs(reinterpret_cast<float4*>(x0 + (size_t)pl * npix) + q, z);
        }
    } else if (vec) {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int pl = 0; pl < nzero; ++pl)
            for (int q = u0 + tid; q < u1; q += kGtThreads) __stcs(reinterpret_cast<float4*>(x0 + (size_t)pl * npix) + q, z);
    } else {
        for (int pl = 0; pl < nzero; ++pl)
            for (int p = u0 + tid; p < u1; p += kGtThreads) x0[(size_t)pl * npix + p] = 0.f;
    }
    double sum = 0.0, sumsq = 0.0;
    if (vec) {
        const float4* g4 = reinterpret_cast<const float4*>(g);
#pragma unroll 2
        for (int q = u0 + tid; q < u1; q += kGtThreads) {
            const float4 a = g4[3 * (size_t)q], c = g4[3 * (size_t)q + 1], e = g4[3 * (size_t)q + 2];
            const float s0 = ((a.x + a.y) + (a.z + a.w)) + ((c.x + c.y) + (c.z + c.w)) + ((e.x + e.y) + (e.z + e.w));
            float s1 = fmaf(a.x, a.x, a.y * a.y) + fmaf(a.z, a.z, a.w * a.w);
            s1 += fmaf(c.x, c.x, c.y * c.y) + fmaf(c.z, c.z, c.w * c.w);
            s1 += fmaf(e.x, e.x, e.y * e.y) + fmaf(e.z, e.z, e.w * e.w);
            sum += (double)s0; sumsq += (double)s1;
        }
    } else {
        for (int p = u0 + tid; p < u1; p += kGtThreads) {
            const float v0 = g[3 * (size_t)p], v1 = g[3 * (size_t)p + 1], v2 = g[3 * (size_t)p + 2];
            sum += (double)(v0 + v1 + v2); sumsq += (double)fmaf(v0, v0, fmaf(v1, v1, v2 * v2));
        }
    }
    sum = warp_sum(sum); sumsq = warp_sum(sumsq);
    if (lane == 0) { s_warp[0][warp] = sum; s_warp[1][warp] = sumsq; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, c = 0.0;
        for (int w = 0; w < kGtThreads / 32; ++w) { a += s_warp[0][w]; c += s_warp[1][w]; }
        s_part[0] = a; s_part[1] = c;
    }
    cluster.sync();
    if (tid == 0) {
        double a = 0.0, c = 0.0;
        for (int r = 0; r < kGtCluster; ++r) {
            const double* rp = cluster.map_shared_rank(s_part, r);
            a += rp[0]; c += rp[1];
        }
        // ref: data_loaders.py:79-82 — population std (ddof 0), no zero guard (Q10): a constant image gives nan / inf
        const double cnt = 3.0 * (double)npix;
        const double mean = a / cnt;
        double var = c / cnt - mean * mean;
        if (var < 0.0) var = 0.0;
        const double inv_std = 1.0 / sqrt(var);
        s_aff[0] = (float)inv_std; s_aff[1] = (float)(-mean * inv_std);
    }
    cluster.sync();     // also keeps every CTA (and its s_part) alive until all ranks have read it
    const float sc = s_aff[0], of = s_aff[1];
    if (vec) {
        const float4* g4 = reinterpret_cast<const float4*>(g);
#pragma unroll 2
        for (int q = u0 + tid; q < u1; q += kGtThreads) {
            const float4 a = __ldcs(g4 + 3 * (size_t)q), c = __ldcs(g4 + 3 * (size_t)q + 1), e = __ldcs(g4 + 3 * (size_t)q + 2);
            __stcs(reinterpret_cast<float4*>(xb) + q,
                   make_float4(fmaf(a.x, sc, of), fmaf(a.w, sc, of), fmaf(c.z, sc, of), fmaf(e.y, sc, of)));
            __stcs(reinterpret_cast<float4*>(xb + npix) + q,
                   make_float4(fmaf(a.y, sc, of), fmaf(c.x, sc, of), fmaf(c.w, sc, of), fmaf(e.z, sc, of)));
            __stcs(reinterpret_cast<float4*>(xb + 2 * (size_t)npix) + q,
                   make_float4(fmaf(a.z, sc, of), fmaf(c.y, sc, of), fmaf(e.x, sc, of), fmaf(e.w, sc, of)));
        }
    } else {
        for (int p = u0 + tid; p < u1; p += kGtThreads)
#pragma unroll
            for (int c = 0; c < 3; ++c) xb[(size_t)c * npix + p] = fmaf(g[3 * (size_t)p + c], sc, of);
    }
    }
    RAC_TL(0);
}

// ---------------------------------------------------------------------------------------------
// gt crop, bulk-copy variant (standalone launches: rac_assemble_input, the end-to-end path of prepare_model_input)
// ---------------------------------------------------------------------------------------------
// Same arithmetic, same cluster-per-image decomposition and the same per-thread summation order as gt_norm_kernel (the
// two are bit-identical), but the bytes move the Blackwell way: every CTA streams its slice of the crop through a
// two-stage shared-memory ring filled by cp.async.bulk (global -> shared, completion on an mbarrier, no registers in
// flight), de-interleaves HWC -> three planes in shared memory, and writes each plane chunk back with cp.async.bulk
// (shared -> global, bulk_group).  One thread issues all copies; the 256 threads only touch shared memory.
// 48 KB of dynamic shared memory per CTA (2 x 12 KB in, 2 x 12 KB out): 4 CTAs per SM, 96 KB of loads in flight per SM.
// Not used next to raster_kernel: four raster CTAs leave ~2 KB of shared memory per SM, which is why the co-resident
// launch of the render path keeps the register-staged kernel above.
constexpr int kGtChunkQuads = 256;                       // one quad (4 px, 48 B in, 3 x 16 B out) per thread per chunk
constexpr int kGtStages = 2;
constexpr int kGtInBytes = kGtChunkQuads * 48;           // 12 KB
constexpr int kGtPlaneBytes = kGtChunkQuads * 16;        //  4 KB per plane chunk
constexpr size_t kGtTmaSmem = (size_t)kGtStages * kGtInBytes + 2 * 3 * kGtPlaneBytes;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy of `bytes` (multiple of 16), completion counted on `bar`
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, const void* src_smem, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(kGtThreads, 4)
gt_norm_tma_kernel(const float* __restrict__ gt_hwc, float* __restrict__ x_nchw, int npix, int C, int ch0, int B) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(128) unsigned char gt_smem[];
    float4* s_in = reinterpret_cast<float4*>(gt_smem);                                        // [kGtStages][kGtChunkQuads * 3]
    float4* s_out = reinterpret_cast<float4*>(gt_smem + (size_t)kGtStages * kGtInBytes);      // [2][3][kGtChunkQuads]
    __shared__ __align__(8) unsigned long long s_full[kGtStages];
    __shared__ double s_warp[2][kGtThreads / 32];
    __shared__ double s_part[2];
    __shared__ float s_aff[2];
    const int rank = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        for (int s = 0; s < kGtStages; ++s) mbar_init(&s_full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    unsigned it = 0;                      // chunks consumed so far by this CTA: stage = it % kGtStages, parity = (it / kGtStages) & 1
    for (int b = blockIdx.y; b < B; b += gridDim.y) {     // cluster-uniform
        const char* g = reinterpret_cast<const char*>(gt_hwc + 3 * (size_t)b * npix);
        float* xb = x_nchw + (size_t)b * C * (size_t)npix + (size_t)ch0 * (size_t)npix;
        const int nunit = npix >> 2;
        const int u0 = (int)((long long)nunit * rank / kGtCluster), u1 = (int)((long long)nunit * (rank + 1) / kGtCluster);
        const int nchunk = (u1 - u0 + kGtChunkQuads - 1) / kGtChunkQuads;
        auto issue = [&](int c, unsigned slot) {          // thread 0: chunk c of this slice -> ring stage `slot`
            const int q0 = u0 + c * kGtChunkQuads;
            const unsigned nq = (unsigned)min(kGtChunkQuads, u1 - q0);
            bulk_load(s_in + (size_t)slot * (kGtChunkQuads * 3), g + 48 * (size_t)q0, nq * 48u, &s_full[slot]);
        };
        double sum = 0.0, sumsq = 0.0;
        // ---- pass 1: statistics ----
        if (tid == 0)
            for (int c = 0; c < min(kGtStages, nchunk); ++c) issue(c, (it + c) % kGtStages);
        for (int c = 0; c < nchunk; ++c, ++it) {
            const unsigned slot = it % kGtStages;
            mbar_wait(&s_full[slot], (it / kGtStages) & 1u);
            const int q = u0 + c * kGtChunkQuads + tid;
            if (q < u1) {
                const float4* p = s_in + (size_t)slot * (kGtChunkQuads * 3) + 3 * tid;
                const float4 a = p[0], cc = p[1], e = p[2];
                const float s0 = ((a.x + a.y) + (a.z + a.w)) + ((cc.x + cc.y) + (cc.z + cc.w)) + ((e.x + e.y) + (e.z + e.w));
                float s1 = fmaf(a.x, a.x, a.y * a.y) + fmaf(a.z, a.z, a.w * a.w);
                s1 += fmaf(cc.x, cc.x, cc.y * cc.y) + fmaf(cc.z, cc.z, cc.w * cc.w);
                s1 += fmaf(e.x, e.x, e.y * e.y) + fmaf(e.z, e.z, e.w * e.w);
                sum += (double)s0; sumsq += (double)s1;
            }
            __syncthreads();              // the stage has been read by everyone
            if (tid == 0 && c + kGtStages < nchunk) issue(c + kGtStages, slot);
        }
        sum = warp_sum(sum); sumsq = warp_sum(sumsq);
        if (lane == 0) { s_warp[0][warp] = sum; s_warp[1][warp] = sumsq; }
        __syncthreads();
        if (tid == 0) {
            double a = 0.0, c = 0.0;
            for (int w = 0; w < kGtThreads / 32; ++w) { a += s_warp[0][w]; c += s_warp[1][w]; }
            s_part[0] = a; s_part[1] = c;
        }
        cluster.sync();
        if (tid == 0) {
            double a = 0.0, c = 0.0;
            for (int r = 0; r < kGtCluster; ++r) {
                const double* rp = cluster.map_shared_rank(s_part, r);
                a += rp[0]; c += rp[1];
            }
            const double cnt = 3.0 * (double)npix;          // ref: data_loaders.py:79-82 (population std, no zero guard)
            const double mean = a / cnt;
            double var = c / cnt - mean * mean;
            if (var < 0.0) var = 0.0;
            const double inv_std = 1.0 / sqrt(var);
            s_aff[0] = (float)inv_std; s_aff[1] = (float)(-mean * inv_std);
        }
        cluster.sync();
        const float sc = s_aff[0], of = s_aff[1];
        // ---- pass 2: normalise + HWC -> CHW (the crop comes back from L2) ----
        if (tid == 0)
            for (int c = 0; c < min(kGtStages, nchunk); ++c) issue(c, (it + c) % kGtStages);
        for (int c = 0; c < nchunk; ++c, ++it) {
            const unsigned slot = it % kGtStages;
            mbar_wait(&s_full[slot], (it / kGtStages) & 1u);
            const int q0 = u0 + c * kGtChunkQuads;
            const int nq = min(kGtChunkQuads, u1 - q0);
            float4* ob = s_out + (size_t)(c & 1) * (3 * kGtChunkQuads);
            if (tid < nq) {
                const float4* p = s_in + (size_t)slot * (kGtChunkQuads * 3) + 3 * tid;
                const float4 a = p[0], cc = p[1], e = p[2];
                ob[tid] = make_float4(fmaf(a.x, sc, of), fmaf(a.w, sc, of), fmaf(cc.z, sc, of), fmaf(e.y, sc, of));
                ob[kGtChunkQuads + tid] = make_float4(fmaf(a.y, sc, of), fmaf(cc.x, sc, of), fmaf(cc.w, sc, of), fmaf(e.z, sc, of));
                ob[2 * kGtChunkQuads + tid] = make_float4(fmaf(a.z, sc, of), fmaf(cc.y, sc, of), fmaf(e.x, sc, of), fmaf(e.w, sc, of));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> visible to the bulk store
            __syncthreads();
            if (tid == 0) {
                for (int pl = 0; pl < 3; ++pl)
                    bulk_store(xb + (size_t)pl * npix + 4 * (size_t)q0, ob + (size_t)pl * kGtChunkQuads, (unsigned)nq * 16u);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                // the OTHER output buffer (chunk c - 1) must have been read before chunk c + 1 overwrites it
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                if (c + kGtStages < nchunk) issue(c + kGtStages, slot);
            }
            __syncthreads();
        }
        if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // all stores complete before the CTA exits
}

// RAC_GT_TMA=0 selects the register-staged kernel for standalone launches too (A/B measurements; default: bulk copies)
static bool gt_tma_enabled() {
    static const int v = [] { const char* e = getenv("RAC_GT_TMA"); return e ? atoi(e) : 1; }();
    return v != 0;
}

// n_clusters: how many images are in flight at once (persistent clusters loop over the batch); <= 0 = one per image
static int launch_gt_norm(const float* d_gt_hwc, float* d_x_nchw, int B, int npix, int use_norm_depth, int zero_render_planes,
                          int n_clusters, cudaStream_t stream, const unsigned* d_live_loads = nullptr, int W = 0, int ntx = 0, int nt = 0) {
    const int C = use_norm_depth ? 7 : 6, ch0 = use_norm_depth ? 4 : 3;
    const int vec = (npix % 4 == 0 && (reinterpret_cast<uintptr_t>(d_gt_hwc) & 15u) == 0 &&
                     (reinterpret_cast<uintptr_t>(d_x_nchw) & 15u) == 0) ? 1 : 0;
    RAC_MAX_CARVEOUT_ONCE(gt_norm_kernel);     // (rac_assemble_input reaches this kernel without render_common)
    // standalone launches (no zero-fill, one cluster per image, nothing co-resident): the bulk-copy variant
    if (vec && !zero_render_planes && n_clusters <= 0 && !d_live_loads && gt_tma_enabled()) {
        static std::atomic<unsigned long long> configured{0};      // bit per device ordinal: function attributes are per device
        int dev = 0;
        RAC_CUDA_CHECK(cudaGetDevice(&dev));
        const unsigned long long bit = 1ull << (dev & 63);
        if (!(configured.load(std::memory_order_acquire) & bit)) {
            RAC_CUDA_CHECK(cudaFuncSetAttribute(gt_norm_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGtTmaSmem));
            configured.fetch_or(bit, std::memory_order_release);
        }
        cudaLaunchConfig_t tc = {};
        tc.gridDim = dim3(kGtCluster, B);
        tc.blockDim = dim3(kGtThreads);
        tc.dynamicSmemBytes = kGtTmaSmem;
        tc.stream = stream;
        cudaLaunchAttribute ta[1];
        ta[0].id = cudaLaunchAttributeClusterDimension;
        ta[0].val.clusterDim.x = kGtCluster; ta[0].val.clusterDim.y = 1; ta[0].val.clusterDim.z = 1;
        tc.attrs = ta; tc.numAttrs = 1;
        RAC_CUDA_CHECK(cudaLaunchKernelEx(&tc, gt_norm_tma_kernel, d_gt_hwc, d_x_nchw, npix, C, ch0, B));
        RAC_LAUNCHED();
        return RAC_OK;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kGtCluster, (n_clusters > 0 && n_clusters < B) ? n_clusters : B);
    cfg.blockDim = dim3(kGtThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kGtCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    RAC_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gt_norm_kernel, d_gt_hwc, d_x_nchw, npix, C, ch0, zero_render_planes ? ch0 : 0, vec, B,
                                       (vec && W > 0 && W % 4 == 0) ? d_live_loads : (const unsigned*)nullptr, W, ntx, nt));
    RAC_LAUNCHED();
    return RAC_OK;
}
