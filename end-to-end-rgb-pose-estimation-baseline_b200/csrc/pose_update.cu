// pose_update.cu — steps 3+4 of the hot path: rotation head -> SO(3), then the SE(3) update.
//
// Replaces rotation_representation.py:4-19 (6D Gram-Schmidt), :22-39 (9D SVD symmetric
// orthogonalisation), :41-118 (axis-angle exp map) and :131-184 (calculate_T_CO_pred),
// forward and backward, as ONE kernel launch each instead of ~25 (+ cuSOLVER gesvdj) launches.
//
// Mapping: one thread per sample, 32 samples per warp.  The per-sample work (a 3x3 Jacobi
// eigen-solve, a few cross products) is a serial dependency chain of a few hundred flops, so
// lanes of a warp cannot usefully share one sample; at B <= 4096 the kernel is launch-latency
// bound either way.  Internals run in fp64 (results are rounded to fp32 once), which keeps the
// SVD head within 1e-6 of the exact polar factor even for ill-conditioned head outputs.
#include "pose_update_dev.cuh"

namespace rac {

template <typename TIN>
__global__ void __launch_bounds__(64)
pose_update_fwd_kernel(const float* __restrict__ y, int rot_repr, const TIN* __restrict__ T_in, double fx, double fy,
                       int B, float* __restrict__ T_out, float* __restrict__ saved) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int hd = head_dim(rot_repr);
    float out[16];
    pose_fwd_sample<TIN>(y + (size_t)b * (hd + 3), rot_repr, T_in + 16 * (size_t)b, fx, fy, out, saved + (size_t)b * RAC_POSE_SAVED_FLOATS);
    float4* o4 = reinterpret_cast<float4*>(T_out + 16 * (size_t)b);
    o4[0] = make_float4(out[0], out[1], out[2], out[3]);
    o4[1] = make_float4(out[4], out[5], out[6], out[7]);
    o4[2] = make_float4(out[8], out[9], out[10], out[11]);
    o4[3] = make_float4(out[12], out[13], out[14], out[15]);
}

__global__ void __launch_bounds__(64)
pose_update_bwd_kernel(const float* __restrict__ gT, const float* __restrict__ saved, int rot_repr, int B, float* __restrict__ gy) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int hd = head_dim(rot_repr);
    pose_bwd_sample(gT + 16 * (size_t)b, saved + (size_t)b * RAC_POSE_SAVED_FLOATS, rot_repr, gy + (size_t)b * (hd + 3));
}

__global__ void __launch_bounds__(64)
rotation_fwd_kernel(const float* __restrict__ in, int rot_repr, float eps, int B, float* __restrict__ R, float* __restrict__ saved) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int hd = head_dim(rot_repr);
    double dR[9];
    head_fwd(rot_repr, in + (size_t)b * hd, (double)eps, dR, saved ? saved + (size_t)b * RAC_POSE_SAVED_FLOATS + kSavedHead : nullptr);
    for (int i = 0; i < 9; ++i) R[9 * (size_t)b + i] = (float)dR[i];
}

__global__ void __launch_bounds__(64)
rotation_bwd_kernel(const float* __restrict__ gR, const float* __restrict__ saved, int rot_repr, int B, float* __restrict__ gin) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int hd = head_dim(rot_repr);
    double G[9];
    for (int i = 0; i < 9; ++i) G[i] = gR[9 * (size_t)b + i];
    head_bwd(rot_repr, saved + (size_t)b * RAC_POSE_SAVED_FLOATS + kSavedHead, G, gin + (size_t)b * hd);
}

}  // namespace rac

using namespace rac;

extern "C" {

int rac_pose_update_fwd(const float* d_y, int rot_repr, const void* d_T_in, int t_in_is_f64, double fx, double fy,
                        int B, float* d_T_out, float* d_saved, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_y && d_T_in && d_T_out && d_saved, "rac_pose_update_fwd: null pointer");
    RAC_REQUIRE(rot_repr >= RAC_ROT_6D && rot_repr <= RAC_ROT_3D, "rac_pose_update_fwd: unknown rot_repr %d", rot_repr);
    RAC_REQUIRE(B > 0, "rac_pose_update_fwd: B=%d", B);
    const int threads = 64, blocks = (B + threads - 1) / threads;
    if (t_in_is_f64) {
        RAC_MAX_CARVEOUT_ONCE(pose_update_fwd_kernel<double>);
        pose_update_fwd_kernel<double><<<blocks, threads, 0, stream>>>(d_y, rot_repr, (const double*)d_T_in, fx, fy, B, d_T_out, d_saved);
    } else {
        RAC_MAX_CARVEOUT_ONCE(pose_update_fwd_kernel<float>);
        pose_update_fwd_kernel<float><<<blocks, threads, 0, stream>>>(d_y, rot_repr, (const float*)d_T_in, fx, fy, B, d_T_out, d_saved);
    }
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_pose_update_bwd(const float* d_gT_out, const float* d_saved, int rot_repr, int B, float* d_gy, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_gT_out && d_saved && d_gy, "rac_pose_update_bwd: null pointer");
    RAC_REQUIRE(rot_repr >= RAC_ROT_6D && rot_repr <= RAC_ROT_3D, "rac_pose_update_bwd: unknown rot_repr %d", rot_repr);
    RAC_REQUIRE(B > 0, "rac_pose_update_bwd: B=%d", B);
    const int threads = 64, blocks = (B + threads - 1) / threads;
    RAC_MAX_CARVEOUT_ONCE(pose_update_bwd_kernel);
    pose_update_bwd_kernel<<<blocks, threads, 0, stream>>>(d_gT_out, d_saved, rot_repr, B, d_gy);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_rotation_fwd(const float* d_in, int rot_repr, float eps, int B, float* d_R, float* d_saved, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_in && d_R, "rac_rotation_fwd: null pointer");
    RAC_REQUIRE(rot_repr >= RAC_ROT_6D && rot_repr <= RAC_ROT_3D, "rac_rotation_fwd: unknown rot_repr %d", rot_repr);
    RAC_REQUIRE(B > 0, "rac_rotation_fwd: B=%d", B);
    const int threads = 64, blocks = (B + threads - 1) / threads;
    rotation_fwd_kernel<<<blocks, threads, 0, stream>>>(d_in, rot_repr, eps, B, d_R, d_saved);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_rotation_bwd(const float* d_gR, const float* d_saved, int rot_repr, int B, float* d_gin, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_gR && d_saved && d_gin, "rac_rotation_bwd: null pointer");
    RAC_REQUIRE(rot_repr >= RAC_ROT_6D && rot_repr <= RAC_ROT_3D, "rac_rotation_bwd: unknown rot_repr %d", rot_repr);
    RAC_REQUIRE(B > 0, "rac_rotation_bwd: B=%d", B);
    const int threads = 64, blocks = (B + threads - 1) / threads;
    rotation_bwd_kernel<<<blocks, threads, 0, stream>>>(d_gR, d_saved, rot_repr, B, d_gin);
    RAC_LAUNCHED();
    return RAC_OK;
}

}  // extern "C"
