// raster_setup.inl — setup_kernel: per-(image, triangle) transform / cull / plane setup + tile binning.
// Included by raster.cu (inside namespace rac).  Spec: RS-V, RS-T in raster.cuh.

__device__ __forceinline__ int floor_div256(int a) { return a >> 8; }             // arithmetic shift
__device__ __forceinline__ int ceil_div256(int a) { return (a + 255) >> 8; }

struct CamPose {
    float r[9];
    float t[3];
    float fx, fy, cx, cy;
};

__device__ __forceinline__ CamPose load_campose(const float* __restrict__ T_CO, const float* __restrict__ K, int b) {
    CamPose c;
    const float* T = T_CO + 16 * (size_t)b;
    c.r[0] = T[0]; c.r[1] = T[1]; c.r[2] = T[2];  c.t[0] = T[3];
    c.r[3] = T[4]; c.r[4] = T[5]; c.r[5] = T[6];  c.t[1] = T[7];
    c.r[6] = T[8]; c.r[7] = T[9]; c.r[8] = T[10]; c.t[2] = T[11];
    const float* k = K + 9 * (size_t)b;
    c.fx = k[0]; c.cx = k[2]; c.fy = k[4]; c.cy = k[5];   // ref: renderer.py:72 (fx, fy, ux, uy only)
    return c;
}

// RS-V. Returns false when the vertex invalidates its triangle.
__device__ __forceinline__ bool project_vertex(const CamPose& c, const float* __restrict__ p, float znear,
                                               int& X, int& Y, float& iz) {
    const float px = p[0], py = p[1], pz = p[2];
    const float xc = __fmaf_rn(c.r[2], pz, __fmaf_rn(c.r[1], py, __fmaf_rn(c.r[0], px, c.t[0])));
    const float yc = __fmaf_rn(c.r[5], pz, __fmaf_rn(c.r[4], py, __fmaf_rn(c.r[3], px, c.t[1])));
    const float zc = __fmaf_rn(c.r[8], pz, __fmaf_rn(c.r[7], py, __fmaf_rn(c.r[6], px, c.t[2])));
    if (!(zc >= znear * kNearRelax)) return false;       // (see RS-V: the near plane itself clips per sample, in depth_key)
    iz = __frcp_rn(zc);      // correctly rounded: == 1.0f / zc
    const float u = __fmaf_rn(c.fx, __fmul_rn(xc, iz), c.cx);
    const float v = __fmaf_rn(c.fy, __fmul_rn(yc, iz), c.cy);
    if (!(fabsf(u) <= kGuardPx) || !(fabsf(v) <= kGuardPx)) return false;
    X = __float2int_rn(__fmul_rn(u, (float)kSubpix));
    Y = __float2int_rn(__fmul_rn(v, (float)kSubpix));
    return true;
}

// Every toucher of a bin adds to the bin's load estimate (binned triangle refs + kLargeWeight per large triangle) with
// a fire-and-forget reduction: a non-zero load marks the bin as non-empty, and its class orders the raster work.
constexpr unsigned kLargeWeight = 16u;
__device__ __forceinline__ void mark_tile(unsigned* loads, unsigned b, unsigned nt, unsigned tile, unsigned weight) {
    atomicAdd(loads + (size_t)b * nt + tile, weight);
}

// work item r of the concatenated bucket lists; s_prefix[k] = number of items in buckets < k (kBuckets + 1 entries)
__device__ __forceinline__ unsigned bucket_entry(const unsigned* s_prefix, const unsigned* __restrict__ buckets, size_t cap, unsigned r) {
    int k = 0;
#pragma unroll
    for (int step = kBuckets / 2; step > 0; step >>= 1)
        if (s_prefix[k + step] <= r) k += step;
    return __ldcg(buckets + (size_t)k * cap + (r - s_prefix[k]));
}

// first warp of a CTA: bucket counts -> exclusive prefix in shared memory (s_prefix[kBuckets] = number of work items)
__device__ __forceinline__ void load_bucket_prefix(const WorkCtl* __restrict__ ctl, unsigned* s_prefix) {
    static_assert(kBuckets <= 32, "one warp scans the bucket counts");
    const int lane = threadIdx.x;
    if (lane < 32) {
        const unsigned c = lane < kBuckets ? __ldcg(&ctl->bucket_cnt[lane]) : 0u;
        unsigned incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane < kBuckets) s_prefix[lane + 1] = incl;
        if (lane == 0) s_prefix[0] = 0u;
    }
}

// One CTA = kSetupThreads (256) consecutive triangles of one image, worked by kSetupCtaThreads (128) threads that own
// TWO triangles each (local ids tid and tid + 128): the two independent load -> project -> cull chains of a thread hide
// each other's latency, the per-CTA fixed work (bin counters, scan, descriptor row, completion) is shared by half as
// many warps, and a barrier waits for 4 warps instead of 8.  Survivors that touch at most 2x2 bins are binned: the CTA
// counts them per bin in shared memory, scans the counts, and publishes
//   desc[b][bin][cta] = (offset << 16) | count      (offset into the CTA's private 1024-B ref region)
//   refs[b][cta][offset + k] = local triangle id (u8) (triangle = cta * 256 + id)
// with plain stores — no global atomics, nothing to zero beforehand.  The others ("large": rare for ModelNet-scale
// meshes, the common case for 12-triangle boxes) are compacted, in triangle order, into the CTA's slice of the image's
// large list: large_tri[b][cta * 256 + k], k < large_cnt[b][cta] — deterministic, unlike an atomic append.
#ifndef RAC_SETUP_CTA_THREADS
#define RAC_SETUP_CTA_THREADS 128
#endif
constexpr int kSetupCtaThreads = RAC_SETUP_CTA_THREADS;
constexpr int kSetupPer = kSetupThreads / kSetupCtaThreads;     // triangles per thread
static_assert(kSetupPer * kSetupCtaThreads == kSetupThreads && kSetupCtaThreads >= 64, "refs copy-out needs >= 64 threads");

__device__ __forceinline__ void
setup_body(const MeshDev* __restrict__ table, const int* __restrict__ mesh_ids, int uniform_mesh,
             const float* __restrict__ T_CO, const float* __restrict__ K, int W, int H, int maxT, TileGrid tg,
             float znear, ImgStats* __restrict__ stats, unsigned* __restrict__ large_cnt, unsigned* __restrict__ large_tri,
             int4* __restrict__ recs, unsigned* __restrict__ desc, unsigned char* __restrict__ refs,
             WorkCtl* __restrict__ ctl, unsigned* __restrict__ loads, unsigned* __restrict__ buckets) {
    __shared__ unsigned s_cnt[kMaxBinTiles];
    __shared__ __align__(16) unsigned char s_refs[4 * kSetupThreads];
    __shared__ unsigned s_total;
    __shared__ unsigned s_off[kMaxBinTiles];
    __shared__ unsigned s_warp[kSetupCtaThreads / 32];
    __shared__ unsigned s_lwarp[kSetupPer][kSetupCtaThreads / 32];
    __shared__ unsigned s_last;
    constexpr int kWarps = kSetupCtaThreads / 32;
    const int b = blockIdx.y;
    const MeshDev& mesh = table[mesh_ids ? mesh_ids[b] : uniform_mesh];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int T = mesh.T;
    if ((int)(blockIdx.x * kSetupThreads) >= T) return;   // CTA beyond this image's mesh: raster reads ceil(T/256) descriptors
    const CamPose cam = load_campose(T_CO, K, b);
    if (tg.binned)
        for (int i = tid; i < tg.nt; i += kSetupCtaThreads) s_cnt[i] = 0;

    bool keep[kSetupPer], fits32[kSetupPer];
    int tx0[kSetupPer], tx1[kSetupPer], ty0[kSetupPer], ty1[kSetupPer];
#pragma unroll
    for (int k = 0; k < kSetupPer; ++k) {
        const int t = blockIdx.x * kSetupThreads + k * kSetupCtaThreads + tid;
        keep[k] = t < T;
        int X0 = 0, Y0 = 0, X1 = 0, Y1 = 0, X2 = 0, Y2 = 0;
        float iz0 = 0.f, iz1 = 0.f, iz2 = 0.f;
        tx0[k] = 0; tx1[k] = -1; ty0[k] = 0; ty1[k] = -1; fits32[k] = false;
        if (keep[k]) {
            const int i0 = mesh.faces[3 * t], i1 = mesh.faces[3 * t + 1], i2 = mesh.faces[3 * t + 2];
            keep[k] = project_vertex(cam, mesh.verts + 3 * (size_t)i0, znear, X0, Y0, iz0);
            keep[k] = keep[k] && project_vertex(cam, mesh.verts + 3 * (size_t)i1, znear, X1, Y1, iz1);
            keep[k] = keep[k] && project_vertex(cam, mesh.verts + 3 * (size_t)i2, znear, X2, Y2, iz2);
        }
        long long area2 = 0;
        bool back = false;         // a back face that is drawn anyway: its part's material is doubleSided (multi-part meshes only)
        if (keep[k]) {
            area2 = (long long)(X1 - X0) * (Y2 - Y0) - (long long)(X2 - X0) * (Y1 - Y0);
            if (area2 > 0 && mesh.parts) back = mesh.parts[mesh.face_part[t]].double_sided != 0;
            keep[k] = area2 < 0 || back;   // RS-T: front faces have negative area in (u, v-down) image coordinates
        }
        if (keep[k]) {
            // pixels that own at least one sample inside the bounding box (sample offsets span 32..224)
            const int xmin = min(X0, min(X1, X2)), xmax = max(X0, max(X1, X2));
            const int ymin = min(Y0, min(Y1, Y2)), ymax = max(Y0, max(Y1, Y2));
            const int px0 = max(ceil_div256(xmin - 224), 0), px1 = min(floor_div256(xmax - 32), W - 1);
            const int py0 = max(ceil_div256(ymin - 224), 0), py1 = min(floor_div256(ymax - 32), H - 1);
            keep[k] = (px0 <= px1) && (py0 <= py1);
            // the UNCLIPPED extent decides whether int32 edge functions are exact: a triangle hanging off an image
            // corner can have an on-screen bounding box of 2x2 bins and still span thousands of pixels
            fits32[k] = (xmax - xmin) <= 64 * kSubpix && (ymax - ymin) <= 64 * kSubpix;
            tx0[k] = px0 / kTileW; tx1[k] = px1 / kTileW; ty0[k] = py0 / kTileH; ty1[k] = py1 / kTileH;
        }
        if (keep[k]) {
            // swap v1 <-> v2 so that the interior is on the non-negative side of all three edges (a drawn back face
            // already has it there)
            if (!back) {
                int sx = X1, sy = Y1; X1 = X2; Y1 = Y2; X2 = sx; Y2 = sy;
                float tz = iz1; iz1 = iz2; iz2 = tz;
            }
            const float areaf = (float)(back ? area2 : -area2);
            const float d1 = __fsub_rn(iz1, iz0), d2 = __fsub_rn(iz2, iz0);
            const float ex1 = (float)(X1 - X0), ey1 = (float)(Y1 - Y0);
            const float ex2 = (float)(X2 - X0), ey2 = (float)(Y2 - Y0);
            const float gx = __fdiv_rn(__fmaf_rn(d1, ey2, -__fmul_rn(d2, ey1)), areaf);
            const float gy = __fdiv_rn(__fmaf_rn(d2, ex1, -__fmul_rn(d1, ex2)), areaf);
            const float* fn = mesh.fnormals + 3 * (size_t)t;
            float nx = cam.r[0] * fn[0] + cam.r[1] * fn[1] + cam.r[2] * fn[2];
            // multi-part meshes: the lowest mantissa bit of nx tells the shader whether the record keeps the mesh's vertex
            // order (1: drawn back face) or the swapped one (0) — it picks the attribute indices accordingly
            if (mesh.parts) nx = __uint_as_float((__float_as_uint(nx) & ~1u) | (back ? 1u : 0u));
            const float ny = cam.r[3] * fn[0] + cam.r[4] * fn[1] + cam.r[5] * fn[2];
            const float nz = cam.r[6] * fn[0] + cam.r[7] * fn[1] + cam.r[8] * fn[2];
            int4* r = recs + 3 * ((size_t)b * maxT + t);
            r[0] = make_int4(X0, Y0, X1, Y1);
            r[1] = make_int4(X2, Y2, __float_as_int(iz0), __float_as_int(gx));
            r[2] = make_int4(__float_as_int(gy), __float_as_int(nx), __float_as_int(ny), __float_as_int(nz));
        }
    }
    // binned: at most 2x2 bins and an unclipped extent of at most 64 px (every edge-function product < 2^28: int32-exact);
    // everything else takes the large list (int64 edge functions)
    bool small[kSetupPer], large[kSetupPer];
#pragma unroll
    for (int k = 0; k < kSetupPer; ++k) {
        small[k] = keep[k] && tg.binned && fits32[k] && (tx1[k] - tx0[k] <= 1) && (ty1[k] - ty0[k] <= 1);
        large[k] = keep[k] && !small[k];
    }
    __syncthreads();   // s_cnt zeroed
    unsigned slot[kSetupPer] = {};           // position within each of the (up to 4) bins, 8 bits each (count <= 256 -> pos <= 255)
#pragma unroll
    for (int k = 0; k < kSetupPer; ++k) {
        if (small[k]) {
            int j = 0;
            for (int ty = ty0[k]; ty <= ty1[k]; ++ty)
                for (int tx = tx0[k]; tx <= tx1[k]; ++tx, ++j) slot[k] |= atomicAdd(&s_cnt[ty * tg.ntx + tx], 1u) << (8 * j);
        }
    }
    // large triangles: CTA-level compaction in triangle order (local ids 0..127 = k 0, 128..255 = k 1)
    unsigned lm[kSetupPer];
#pragma unroll
    for (int k = 0; k < kSetupPer; ++k) {
        lm[k] = __ballot_sync(0xffffffffu, large[k]);
        if (lane == 0) s_lwarp[k][warp] = (unsigned)__popc(lm[k]);
    }
    __syncthreads();   // (also: every counting atomic above is done, s_cnt is complete)
    {
        unsigned total = 0, base[kSetupPer];
#pragma unroll
        for (int k = 0; k < kSetupPer; ++k) {
            base[k] = total;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const unsigned v = s_lwarp[k][w];
                if (w < warp) base[k] += v;
                total += v;
            }
        }
        if (tid == 0) {
            large_cnt[(size_t)b * tg.ncta_max + blockIdx.x] = total;
            if (total) atomicAdd(&stats[b].nlarge, total);
        }
#pragma unroll
        for (int k = 0; k < kSetupPer; ++k) {
            if (large[k]) {
                const int t = blockIdx.x * kSetupThreads + k * kSetupCtaThreads + tid;
                large_tri[(size_t)b * maxT + (size_t)blockIdx.x * kSetupThreads + base[k] + __popc(lm[k] & ((1u << lane) - 1u))] = (unsigned)t;
                for (int ty = ty0[k]; ty <= ty1[k]; ++ty)
                    for (int tx = tx0[k]; tx <= tx1[k]; ++tx) mark_tile(loads, (unsigned)b, (unsigned)tg.nt, (unsigned)(ty * tg.ntx + tx), kLargeWeight);
            }
        }
    }
    if (tg.binned) {   // CTA-uniform
        // exclusive scan of s_cnt[0..nt) -> s_off: thread i owns items [i*per, (i+1)*per)
        const int per = (tg.nt + kSetupCtaThreads - 1) / kSetupCtaThreads;
        const int i_lo = min(tid * per, tg.nt), i_hi = min(i_lo + per, tg.nt);
        unsigned mine = 0;
        for (int i = i_lo; i < i_hi; ++i) mine += s_cnt[i];
        unsigned incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        unsigned run = incl - mine;
        for (int w = 0; w < warp; ++w) run += s_warp[w];
        unsigned* dimg = desc + (size_t)b * tg.nt * tg.ncta_max + blockIdx.x;
        for (int i = i_lo; i < i_hi; ++i) {
            const unsigned c = s_cnt[i];
            s_off[i] = run;
            dimg[(size_t)i * tg.ncta_max] = (run << 16) | c;    // offset <= 1024, count <= 256
            if (c) mark_tile(loads, (unsigned)b, (unsigned)tg.nt, (unsigned)i, c);
            run += c;
        }
        if (tid == kSetupCtaThreads - 1) s_total = run;       // this thread owns the last bins: run = all refs of the CTA
        __syncthreads();
        // refs: grouped by bin in shared memory first, then copied out with 16-B stores (not ~350 scattered 1-B ones)
#pragma unroll
        for (int k = 0; k < kSetupPer; ++k) {
            if (small[k]) {
                int j = 0;
                for (int ty = ty0[k]; ty <= ty1[k]; ++ty)
                    for (int tx = tx0[k]; tx <= tx1[k]; ++tx, ++j)
                        s_refs[s_off[ty * tg.ntx + tx] + ((slot[k] >> (8 * j)) & 0xFFu)] = (unsigned char)(k * kSetupCtaThreads + tid);
            }
        }
        __syncthreads();
        {
            uint4* region = reinterpret_cast<uint4*>(refs + (((size_t)b * tg.ncta_max + blockIdx.x) << 10));
            const int n16 = ((int)s_total + 15) >> 4;
            if (tid < n16) region[tid] = reinterpret_cast<const uint4*>(s_refs)[tid];     // <= 64 stores of 16 B
        }
    }
    // ---- image completion: the CTA that finishes an image last publishes the image's non-empty bins by load class ----
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(&stats[b].ndone, 1u) == (unsigned)((T + kSetupThreads - 1) / kSetupThreads) - 1u) ? 1u : 0u;
    __syncthreads();
    if (s_last) {      // CTA-uniform
        __threadfence();
        const size_t cap = (size_t)gridDim.y * tg.nt;
        for (int i = tid; i < tg.nt; i += kSetupCtaThreads) {
            const unsigned ld = __ldcg(loads + (size_t)b * tg.nt + i);
            if (ld) {
                const unsigned k = (unsigned)(kBuckets - 1) - min(ld >> kBucketShift, (unsigned)(kBuckets - 1));
                buckets[k * cap + atomicAdd(&ctl->bucket_cnt[k], 1u)] = (unsigned)b * (unsigned)tg.nt + (unsigned)i;
            }
        }
    }
}

#ifndef RAC_SETUP_MIN_CTAS
#define RAC_SETUP_MIN_CTAS 5
#endif
__global__ void __launch_bounds__(kSetupCtaThreads, RAC_SETUP_MIN_CTAS)
setup_kernel(const MeshDev* __restrict__ table, const int* __restrict__ mesh_ids, int uniform_mesh,
             const float* __restrict__ T_CO, const float* __restrict__ K, int W, int H, int maxT, TileGrid tg,
             float znear, ImgStats* __restrict__ stats, unsigned* __restrict__ large_cnt, unsigned* __restrict__ large_tri,
             int4* __restrict__ recs, unsigned* __restrict__ desc, unsigned char* __restrict__ refs,
             WorkCtl* __restrict__ ctl, unsigned* __restrict__ loads, unsigned* __restrict__ buckets) {
    RAC_TL(1);
    setup_body(table, mesh_ids, uniform_mesh, T_CO, K, W, H, maxT, tg, znear, stats, large_cnt, large_tri, recs, desc, refs, ctl,
               loads, buckets);
    RAC_TL(1);
}
