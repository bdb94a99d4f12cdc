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
    double gmean, gstd;
};

__device__ __forceinline__ NormConsts load_norm(const ImgStats& st, size_t npix, bool want_depth_norm, bool have_gt) {
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
    n.gmean = 0.0; n.gstd = 1.0;
    if (have_gt) {
        // ref: data_loaders.py:79-82 — population std, no zero guard (Q10): a constant image gives inf/nan
        const double cnt = 3.0 * (double)npix;
        n.gmean = st.gsum / cnt;
        double var = st.gsumsq / cnt - n.gmean * n.gmean;
        if (var < 0.0) var = 0.0;
        n.gstd = sqrt(var);
    }
    return n;
}

__device__ __forceinline__ float norm_depth1(float d, const NormConsts& n) {
    if (!n.depth_norm) return d;
    return d > 0.01f ? __fdiv_rn(__fsub_rn(d, n.dmean), n.dstd) : 0.0f;
}
__device__ __forceinline__ float norm_gt1(float g, const NormConsts& n) {
    return (float)(((double)g - n.gmean) / n.gstd);
}
__device__ __forceinline__ float norm_col1(unsigned q, const NormConsts& n) {
    return n.inv_qmax_valid != 0.0f ? __fdiv_rn((float)q, n.qmax) : 0.0f;
}

// Tile flags written by setup_kernel: 0 = no triangle touched the tile -> background (the raster kernel never wrote
// its RGBA8 / depth there).  flags == nullptr: every pixel is valid (prepare_model_input on an existing render).
struct FlagView {
    const unsigned* flags;   // [B][nt]
    int W, ntx, nt;
};
__device__ __forceinline__ bool pixel_live(const FlagView& fv, int b, int p) {
    if (!fv.flags) return true;
    const int y = p / fv.W, x = p - y * fv.W;
    return fv.flags[(size_t)b * fv.nt + (y / kTileH) * fv.ntx + (x / kTileW)] != 0u;
}

// Everything a finalize pass needs.  mode 0: NCHW concat (prepare_model_input), mode 1: render_batch outputs.
struct FinArgs {
    const ImgStats* stats;
    const unsigned* rgba;      // resolved RGBA8 framebuffer (workspace)
    const float* init_hwc;     // SRC_F32 only: an existing f32 HWC render
    const float* depth;        // metric depth (workspace or the caller's buffer)
    const float* gt_hwc;
    int npix, use_norm_depth;
    float* x_nchw;
    float* rgb_hwc_out;        // mode 0: optional HWC copy of the normalised render; mode 1: the colour output
    unsigned* rgba_copy;       // mode 1: optional RGBA8 copy
    float* depth_bg;           // caller's depth buffer whose background tiles still need zeroing (or nullptr)
    FlagView fv;
    int mode, vec;
};

// ImgStats through L2 (ld.global.cg): inside the fused kernel they were produced by other SMs' atomics, and this
// SM's L1 may hold a stale line from an earlier plain read.
__device__ __forceinline__ ImgStats load_stats_cg(const ImgStats* p) {
    ImgStats st;
    const int4* src = reinterpret_cast<const int4*>(p);
    int4* dst = reinterpret_cast<int4*>(&st);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = __ldcg(src + i);
    return st;
}

// One thread = 4 consecutive pixels of one image (requires W*H % 4 == 0 and 16-B aligned planes).
// SRC_F32: the render comes as f32 HWC (prepare_model_input on an existing render) instead of RGBA8.
template <bool SRC_F32>
__device__ __forceinline__ void fin_concat_quad(const FinArgs& A, const NormConsts& n, int b, int q) {
    const unsigned* __restrict__ rgba = A.rgba;
    const float* __restrict__ init_hwc = A.init_hwc;
    const float* __restrict__ depth = A.depth;
    const float* __restrict__ gt_hwc = A.gt_hwc;
    float* __restrict__ x_nchw = A.x_nchw;
    float* __restrict__ rgb_hwc_out = A.rgb_hwc_out;
    float* __restrict__ depth_bg = A.depth_bg;
    const FlagView& fv = A.fv;
    const int npix = A.npix, use_norm_depth = A.use_norm_depth;
    const size_t pbase = (size_t)b * npix + 4 * (size_t)q;
    const int C = use_norm_depth ? 7 : 6;
    float* xb = x_nchw + (size_t)b * C * npix + 4 * (size_t)q;
    int ch = 0;
    // W % 4 == 0: the quad lies in one row and one tile; otherwise test each pixel
    bool live[4];
    live[0] = pixel_live(fv, b, 4 * q);
    if ((fv.W & 3) == 0 || !fv.flags) { live[1] = live[2] = live[3] = live[0]; }
    else { live[1] = pixel_live(fv, b, 4 * q + 1); live[2] = pixel_live(fv, b, 4 * q + 2); live[3] = pixel_live(fv, b, 4 * q + 3); }
    const bool any_live = live[0] || live[1] || live[2] || live[3];
    const bool all_live = live[0] && live[1] && live[2] && live[3];
    if (depth_bg && !all_live) {   // the caller's depth buffer: background tiles were never written by the raster kernel
#pragma unroll
        for (int i = 0; i < 4; ++i) if (!live[i]) depth_bg[pbase + i] = 0.0f;
    }
    if (use_norm_depth) {
        float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
        if (any_live) {
            d = __ldcg(reinterpret_cast<const float4*>(depth + pbase));
            if (!live[0]) d.x = 0.f; if (!live[1]) d.y = 0.f; if (!live[2]) d.z = 0.f; if (!live[3]) d.w = 0.f;
        }
        float4 o;
        o.x = norm_depth1(d.x, n); o.y = norm_depth1(d.y, n); o.z = norm_depth1(d.z, n); o.w = norm_depth1(d.w, n);
        __stcs(reinterpret_cast<float4*>(xb), o);
        ch = 1;
    }
    float r[4], g[4], bl[4];
    if (SRC_F32) {
        const float4* s = reinterpret_cast<const float4*>(init_hwc + 3 * pbase);
        const float4 a = s[0], c = s[1], e = s[2];
        r[0] = a.x; g[0] = a.y; bl[0] = a.z; r[1] = a.w; g[1] = c.x; bl[1] = c.y;
        r[2] = c.z; g[2] = c.w; bl[2] = e.x; r[3] = e.y; g[3] = e.z; bl[3] = e.w;
    } else {
        uint4 c = make_uint4(0u, 0u, 0u, 0u);
        if (any_live) c = __ldcg(reinterpret_cast<const uint4*>(rgba + pbase));
        const unsigned cc[4] = {live[0] ? c.x : 0u, live[1] ? c.y : 0u, live[2] ? c.z : 0u, live[3] ? c.w : 0u};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            r[i] = norm_col1(cc[i] & 0xFFu, n); g[i] = norm_col1((cc[i] >> 8) & 0xFFu, n);
            bl[i] = norm_col1((cc[i] >> 16) & 0xFFu, n);
        }
    }
    __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 0) * npix), make_float4(r[0], r[1], r[2], r[3]));
    __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 1) * npix), make_float4(g[0], g[1], g[2], g[3]));
    __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 2) * npix), make_float4(bl[0], bl[1], bl[2], bl[3]));
    if (rgb_hwc_out) {
        float4* o = reinterpret_cast<float4*>(rgb_hwc_out + 3 * pbase);
        __stcs(o + 0, make_float4(r[0], g[0], bl[0], r[1]));
        __stcs(o + 1, make_float4(g[1], bl[1], r[2], g[2]));
        __stcs(o + 2, make_float4(bl[2], r[3], g[3], bl[3]));
    }
    {
        const float4* s = reinterpret_cast<const float4*>(gt_hwc + 3 * pbase);
        const float4 a = __ldcs(s), c = __ldcs(s + 1), e = __ldcs(s + 2);
        const float gr[4] = {a.x, a.w, c.z, e.y}, gg[4] = {a.y, c.x, c.w, e.z}, gb[4] = {a.z, c.y, e.x, e.w};
        __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 3) * npix),
               make_float4(norm_gt1(gr[0], n), norm_gt1(gr[1], n), norm_gt1(gr[2], n), norm_gt1(gr[3], n)));
        __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 4) * npix),
               make_float4(norm_gt1(gg[0], n), norm_gt1(gg[1], n), norm_gt1(gg[2], n), norm_gt1(gg[3], n)));
        __stcs(reinterpret_cast<float4*>(xb + (size_t)(ch + 5) * npix),
               make_float4(norm_gt1(gb[0], n), norm_gt1(gb[1], n), norm_gt1(gb[2], n), norm_gt1(gb[3], n)));
    }
}

// scalar fallback for shapes the vector kernel cannot take (W*H % 4 != 0 or unaligned pointers)
template <bool SRC_F32>
__device__ __forceinline__ void fin_concat_px(const FinArgs& A, const NormConsts& n, int b, int p) {
    const unsigned* __restrict__ rgba = A.rgba;
    const float* __restrict__ init_hwc = A.init_hwc;
    const float* __restrict__ depth = A.depth;
    const float* __restrict__ gt_hwc = A.gt_hwc;
    float* __restrict__ x_nchw = A.x_nchw;
    float* __restrict__ rgb_hwc_out = A.rgb_hwc_out;
    float* __restrict__ depth_bg = A.depth_bg;
    const FlagView& fv = A.fv;
    const int npix = A.npix, use_norm_depth = A.use_norm_depth;
    const size_t pi = (size_t)b * npix + p;
    const int C = use_norm_depth ? 7 : 6;
    float* xb = x_nchw + (size_t)b * C * npix + p;
    int ch = 0;
    const bool live = pixel_live(fv, b, p);
    if (depth_bg && !live) depth_bg[pi] = 0.0f;
    if (use_norm_depth) { xb[0] = norm_depth1(live ? __ldcg(depth + pi) : 0.0f, n); ch = 1; }
    float r, g, bl;
    if (SRC_F32) { r = init_hwc[3 * pi]; g = init_hwc[3 * pi + 1]; bl = init_hwc[3 * pi + 2]; }
    else {
        const unsigned c = live ? __ldcg(rgba + pi) : 0u;
        r = norm_col1(c & 0xFFu, n); g = norm_col1((c >> 8) & 0xFFu, n); bl = norm_col1((c >> 16) & 0xFFu, n);
    }
    xb[(size_t)(ch + 0) * npix] = r; xb[(size_t)(ch + 1) * npix] = g; xb[(size_t)(ch + 2) * npix] = bl;
    if (rgb_hwc_out) { rgb_hwc_out[3 * pi] = r; rgb_hwc_out[3 * pi + 1] = g; rgb_hwc_out[3 * pi + 2] = bl; }
    xb[(size_t)(ch + 3) * npix] = norm_gt1(gt_hwc[3 * pi], n);
    xb[(size_t)(ch + 4) * npix] = norm_gt1(gt_hwc[3 * pi + 1], n);
    xb[(size_t)(ch + 5) * npix] = norm_gt1(gt_hwc[3 * pi + 2], n);
}

// render_batch API mode: f32 HWC colour (+ optional copy of the RGBA8 framebuffer)
__device__ __forceinline__ void fin_render_px(const FinArgs& A, const NormConsts& n, int b, int p) {
    float* __restrict__ rgb_hwc = A.rgb_hwc_out;
    unsigned* __restrict__ rgba_copy = A.rgba_copy;
    float* __restrict__ depth_bg = A.depth_bg;
    const size_t pi = (size_t)b * A.npix + p;
    const bool live = pixel_live(A.fv, b, p);
    if (depth_bg && !live) depth_bg[pi] = 0.0f;
    const unsigned c = live ? __ldcg(A.rgba + pi) : 0u;
    if (rgba_copy) rgba_copy[pi] = c;
    if (rgb_hwc) {
        rgb_hwc[3 * pi] = norm_col1(c & 0xFFu, n);
        rgb_hwc[3 * pi + 1] = norm_col1((c >> 8) & 0xFFu, n);
        rgb_hwc[3 * pi + 2] = norm_col1((c >> 16) & 0xFFu, n);
    }
}

// One finalize work item = kFinChunkPix consecutive pixels of image b, done by one CTA (fused into raster_kernel once
// the image's last tile has been rasterised; also what the standalone kernel below runs per CTA).
constexpr int kFinChunkPix = 4096;

template <bool SRC_F32>
__device__ __forceinline__ void finalize_item(const FinArgs& A, int b, int chunk) {
    const ImgStats st = load_stats_cg(A.stats + b);
    const NormConsts n = load_norm(st, (size_t)A.npix, A.use_norm_depth != 0, A.mode == 0);
    const int p0 = chunk * kFinChunkPix, p1 = min(A.npix, p0 + kFinChunkPix);
    if (A.mode == 0) {
        if (A.vec) {
            for (int q = (p0 >> 2) + (int)threadIdx.x; 4 * q < p1; q += (int)blockDim.x) fin_concat_quad<SRC_F32>(A, n, b, q);
        } else {
            for (int p = p0 + (int)threadIdx.x; p < p1; p += (int)blockDim.x) fin_concat_px<SRC_F32>(A, n, b, p);
        }
    } else {
        for (int p = p0 + (int)threadIdx.x; p < p1; p += (int)blockDim.x) fin_render_px(A, n, b, p);
    }
}

// standalone finalize: grid (chunks, B).  Used by rac_assemble_input (SRC_F32) — the render paths run the same
// items inside raster_kernel.
template <bool SRC_F32>
__global__ void __launch_bounds__(256)
finalize_kernel(FinArgs A) {
    finalize_item<SRC_F32>(A, blockIdx.y, blockIdx.x);
}

// prepare_model_input alone: per-image gt (and depth) statistics
__global__ void __launch_bounds__(256)
assemble_stats_kernel(ImgStats* __restrict__ stats, const float* __restrict__ gt_hwc,
                      const float* __restrict__ depth, int npix, int chunks) {
    const int b = blockIdx.y;
    const int per = (npix + chunks - 1) / chunks;
    const int p0 = blockIdx.x * per, p1 = min(p0 + per, npix);
    double gs = 0, gss = 0, ds = 0, dss = 0;
    unsigned dc = 0;
    float dmax = 0.f;
    {
        // 16-B aligned body with 128-bit streaming loads, scalar head / tail
        const float* g = gt_hwc + 3 * ((size_t)b * npix + p0);
        const int n = 3 * (p1 - p0);
        int head = (int)(((16u - (unsigned)(reinterpret_cast<uintptr_t>(g) & 15u)) & 15u) >> 2);
        if (head > n) head = n;
        const int nv = (n - head) >> 2;
        const float4* g4 = reinterpret_cast<const float4*>(g + head);
        for (int i = threadIdx.x; i < nv; i += blockDim.x) {
            const float4 v = __ldg(g4 + i);
            gs += ((double)v.x + (double)v.y) + ((double)v.z + (double)v.w);
            gss += ((double)v.x * v.x + (double)v.y * v.y) + ((double)v.z * v.z + (double)v.w * v.w);
        }
        for (int i = threadIdx.x; i < head; i += blockDim.x) { const double v = (double)g[i]; gs += v; gss += v * v; }
        for (int i = head + 4 * nv + threadIdx.x; i < n; i += blockDim.x) { const double v = (double)g[i]; gs += v; gss += v * v; }
    }
    if (depth) {
        const float* d = depth + (size_t)b * npix;
        for (int i = p0 + threadIdx.x; i < p1; i += blockDim.x) {
            const float v = d[i];
            dmax = fmaxf(dmax, v);
            if (v > 0.01f) { dc++; ds += (double)v; dss += (double)v * v; }
        }
    }
    __shared__ double rd[4][8];
    __shared__ unsigned ru[2][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    gs = warp_sum(gs); gss = warp_sum(gss); ds = warp_sum(ds); dss = warp_sum(dss);
    dc = (unsigned)__reduce_add_sync(0xffffffffu, dc);
    const unsigned dm = warp_max(__float_as_uint(fmaxf(dmax, 0.0f)));
    if (lane == 0) { rd[0][warp] = gs; rd[1][warp] = gss; rd[2][warp] = ds; rd[3][warp] = dss; ru[0][warp] = dc; ru[1][warp] = dm; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, c = 0, e = 0, f = 0; unsigned u = 0, m = 0;
        for (int w = 0; w < (int)blockDim.x / 32; ++w) { a += rd[0][w]; c += rd[1][w]; e += rd[2][w]; f += rd[3][w]; u += ru[0][w]; m = max(m, ru[1][w]); }
        ImgStats* st = stats + b;
        atomicAdd(&st->gsum, a); atomicAdd(&st->gsumsq, c);
        if (u) { atomicAdd(&st->dcount, u); atomicAdd(&st->dsum, e); atomicAdd(&st->dsumsq, f); }
        if (m) atomicMax(&st->dmax_bits, m);
    }
}
