// add_loss.cu — step 5 of the hot path: fused transform-and-reduce point-matching losses.
//
// Replaces loss.py:5-18 (ADD-L1), :20-33 (ADD-L2), :60-78 (transform_pts), :81-97
// (disentangled ADD-L1) and :101-122 (scaled disentangled), forward and backward.
// The reference materialises 2-8 (B,N,3) transformed point clouds per loss and reduces
// them with separate kernels; here one CTA per sample streams the N points once, keeps
// the |.|-sums AND the sign-sums that make up d(loss)/dT in registers, block-reduces,
// and the last CTA to finish folds the per-sample losses into the batch mean in a fixed
// order (deterministic, no float atomics).  The backward is then a 16-float scale per sample.
#include "add_loss_dev.cuh"

namespace rac {

// accumulators: [0] loss numerator, [1..9] S_ij = sum sign(d_i) p_j (rotation part),
// [10..12] s_i = sum sign(d_i) (translation part), [13] scaled-mode denominator numerator,
// [14..16] spare
__global__ void __launch_bounds__(kLossThreads)
add_loss_fwd_kernel(const float* __restrict__ T_a, const float* __restrict__ T_b, const float* __restrict__ T_prev,
                    const float* __restrict__ pts, int B, int N, int mode,
                    float* __restrict__ per_sample, float* __restrict__ mean_out,
                    float* __restrict__ grad_a, float* __restrict__ grad_b, unsigned* __restrict__ counter) {
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const Pose34 A = load_pose(T_a + 16 * (size_t)b);
    const Pose34 Bp = load_pose(T_b + 16 * (size_t)b);
    Pose34 P = A;
    if (mode == RAC_LOSS_DISENT_SCALED) P = load_pose(T_prev + 16 * (size_t)b);
    float acc[kNAcc];
#pragma unroll
    for (int i = 0; i < kNAcc; ++i) acc[i] = 0.f;

    const float* p = pts + 3 * (size_t)b * N;
    for (int n = tid; n < N; n += kLossThreads) {
        const float x = p[3 * n], y = p[3 * n + 1], z = p[3 * n + 2];
        loss_point(mode, A, Bp, P, x, y, z, acc);
    }
    __shared__ float red[kNAcc][kLossThreads / 32];
    __shared__ bool is_last;
#pragma unroll
    for (int i = 0; i < 14; ++i) {
        const float v = warp_sum(acc[i]);
        if (lane == 0) red[i][warp] = v;
    }
    __syncthreads();
    if (tid == 0) {
        float tot[14];
        for (int i = 0; i < 14; ++i) {
            float s = 0.f;
            for (int w = 0; w < kLossThreads / 32; ++w) s += red[i][w];
            tot[i] = s;
        }
        float loss, ga[16], gb[16];
        loss_finish(mode, tot, N, loss, ga, gb);
        per_sample[b] = loss;
        {
            if (grad_a) for (int i = 0; i < 16; ++i) grad_a[16 * (size_t)b + i] = ga[i];
            if (grad_b) for (int i = 0; i < 16; ++i) grad_b[16 * (size_t)b + i] = gb[i];
        }
        __threadfence();
        const unsigned done = atomicAdd(counter, 1u);
        is_last = (done == (unsigned)B - 1u);
    }
    __syncthreads();
    if (is_last && warp == 0) {
        __threadfence();
        float s = 0.f;
        for (int i = lane; i < B; i += 32) s += __ldcg(per_sample + i);
        s = warp_sum(s);
        if (lane == 0) {
            if (mean_out) *mean_out = s / (float)B;
            *counter = 0u;
        }
    }
}

__global__ void __launch_bounds__(256)
add_loss_bwd_kernel(const float* __restrict__ grad, const float* __restrict__ w, int n16, float* __restrict__ gT) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n16) return;
    gT[i] = grad[i] * w[i >> 4];
}

// ---- transform_pts ------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
transform_pts_fwd_kernel(const float* __restrict__ T, const float* __restrict__ pts, int B, int M, int N, float* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t total = (size_t)B * M * N;
    if (i >= total) return;
    const int n = (int)(i % N);
    const size_t bm = i / N;
    const int b = (int)(bm / M);
    const Pose34 P = load_pose(T + 16 * bm);
    const float* p = pts + 3 * ((size_t)b * N + n);
    float o[3];
    xform(P.r, P.t, p[0], p[1], p[2], o);
    out[3 * i] = o[0]; out[3 * i + 1] = o[1]; out[3 * i + 2] = o[2];
}

// gT: one CTA per (b,m) reduces over N
__global__ void __launch_bounds__(kLossThreads)
transform_pts_bwd_T_kernel(const float* __restrict__ gout, const float* __restrict__ pts, int M, int N, float* __restrict__ gT) {
    const size_t bm = blockIdx.x;
    const int b = (int)(bm / M);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float acc[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = 0.f;
    const float* g = gout + 3 * bm * N;
    const float* p = pts + 3 * (size_t)b * N;
    for (int n = tid; n < N; n += kLossThreads) {
        const float x = p[3 * n], y = p[3 * n + 1], z = p[3 * n + 2];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float gi = g[3 * n + i];
            acc[4 * i] += gi * x; acc[4 * i + 1] += gi * y; acc[4 * i + 2] += gi * z; acc[4 * i + 3] += gi;
        }
    }
    __shared__ float red[12][kLossThreads / 32];
#pragma unroll
    for (int i = 0; i < 12; ++i) {
        const float v = warp_sum(acc[i]);
        if (lane == 0) red[i][warp] = v;
    }
    __syncthreads();
    if (tid < 16) {
        float s = 0.f;
        if (tid < 12) for (int w = 0; w < kLossThreads / 32; ++w) s += red[tid][w];
        gT[16 * bm + tid] = s;
    }
}

__global__ void __launch_bounds__(256)
transform_pts_bwd_pts_kernel(const float* __restrict__ gout, const float* __restrict__ T, int B, int M, int N, float* __restrict__ gpts) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)B * N) return;
    const int n = (int)(i % N);
    const int b = (int)(i / N);
    float gx = 0.f, gy = 0.f, gz = 0.f;
    for (int m = 0; m < M; ++m) {
        const size_t bm = (size_t)b * M + m;
        const float* t = T + 16 * bm;
        const float* g = gout + 3 * (bm * N + n);
        gx += t[0] * g[0] + t[4] * g[1] + t[8] * g[2];
        gy += t[1] * g[0] + t[5] * g[1] + t[9] * g[2];
        gz += t[2] * g[0] + t[6] * g[1] + t[10] * g[2];
    }
    gpts[3 * i] = gx; gpts[3 * i + 1] = gy; gpts[3 * i + 2] = gz;
}

}  // namespace rac

using namespace rac;

extern "C" {

int rac_add_loss_fwd(const float* d_T_a, const float* d_T_b, const float* d_T_prev, const float* d_pts, int B, int N,
                     int mode, float* d_per_sample, float* d_mean, float* d_grad_a, float* d_grad_b,
                     unsigned int* d_counter, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_T_a && d_T_b && d_pts && d_per_sample && d_counter, "rac_add_loss_fwd: null pointer");
    RAC_REQUIRE(mode >= RAC_LOSS_ADD_L1 && mode <= RAC_LOSS_DISENT_SCALED, "rac_add_loss_fwd: unknown mode %d", mode);
    RAC_REQUIRE(mode != RAC_LOSS_DISENT_SCALED || d_T_prev, "rac_add_loss_fwd: scaled mode needs T_prev");
    RAC_REQUIRE(B > 0 && N > 0, "rac_add_loss_fwd: B=%d N=%d", B, N);
    RAC_MAX_CARVEOUT_ONCE(add_loss_fwd_kernel);
    add_loss_fwd_kernel<<<B, kLossThreads, 0, stream>>>(d_T_a, d_T_b, d_T_prev, d_pts, B, N, mode, d_per_sample, d_mean,
                                                        d_grad_a, d_grad_b, d_counter);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_add_loss_bwd(const float* d_grad, const float* d_w, int B, float* d_gT, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_grad && d_w && d_gT, "rac_add_loss_bwd: null pointer");
    RAC_REQUIRE(B > 0, "rac_add_loss_bwd: B=%d", B);
    const int n16 = 16 * B;
    RAC_MAX_CARVEOUT_ONCE(add_loss_bwd_kernel);
    add_loss_bwd_kernel<<<(n16 + 255) / 256, 256, 0, stream>>>(d_grad, d_w, n16, d_gT);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_transform_pts_fwd(const float* d_T, const float* d_pts, int B, int M, int N, float* d_out, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_T && d_pts && d_out, "rac_transform_pts_fwd: null pointer");
    RAC_REQUIRE(B > 0 && M > 0 && N > 0, "rac_transform_pts_fwd: B=%d M=%d N=%d", B, M, N);
    const size_t total = (size_t)B * M * N;
    transform_pts_fwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(d_T, d_pts, B, M, N, d_out);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_transform_pts_bwd(const float* d_gout, const float* d_T, const float* d_pts, int B, int M, int N,
                          float* d_gT, float* d_gpts, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_gout && d_T && d_pts, "rac_transform_pts_bwd: null pointer");
    RAC_REQUIRE(B > 0 && M > 0 && N > 0, "rac_transform_pts_bwd: B=%d M=%d N=%d", B, M, N);
    if (d_gT) {
        transform_pts_bwd_T_kernel<<<(unsigned)((size_t)B * M), kLossThreads, 0, stream>>>(d_gout, d_pts, M, N, d_gT);
        RAC_LAUNCHED();
    }
    if (d_gpts) {
        const size_t total = (size_t)B * N;
        transform_pts_bwd_pts_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(d_gout, d_T, B, M, N, d_gpts);
        RAC_LAUNCHED();
    }
    return RAC_OK;
}

}  // extern "C"
