// pose_loss_step.cu — the whole post-CNN half of one refinement iteration in ONE launch.
//
// ref: train_model.py:245-247 + :276-285 / test_model.py:101-103 — calculate_T_CO_pred (rotation_representation.py:131-184),
// the point-matching loss (loss.py:5-33,81-122), its batch mean and `.backward()` down to the CNN's output.  The
// reference-named Python functions keep their own kernels (pose_update.cu, add_loss.cu: four launches behind two
// autograd.Functions, because the unchanged drivers call them one by one); a caller that owns its loop
// (rac.RenderCompareStep, bench.py's serial chain) takes this entry point: one CTA per sample computes T_new in its
// first thread, streams the sample's N points once for the loss and d(loss)/dT_new, applies the 1/B of the batch mean
// and runs the pose update's backward — no intermediate tensor, no autograd graph, one launch instead of four plus
// three copies.  Per-sample arithmetic is the same device code the stand-alone kernels run (pose_update_dev.cuh,
// add_loss_dev.cuh): results are bit-identical to the four-launch path.
#include "pose_update_dev.cuh"
#include "add_loss_dev.cuh"

namespace rac {

template <typename TIN>
__global__ void __launch_bounds__(kLossThreads)
pose_loss_step_kernel(const float* __restrict__ y, int rot_repr, const TIN* T_in, double fx, double fy,
                      const float* __restrict__ T_gt, const float* T_prev, const float* __restrict__ pts,
                      int B, int N, int mode, int pred_is_a, float* __restrict__ T_new, float* __restrict__ per_sample,
                      float* __restrict__ mean_out, float* __restrict__ grad_y, float* T_feedback, unsigned* __restrict__ counter) {
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int D = head_dim(rot_repr) + 3;
    __shared__ __align__(16) float s_T[16];
    __shared__ float s_sv[RAC_POSE_SAVED_FLOATS];
    if (tid == 0) {
        pose_fwd_sample<TIN>(y + (size_t)b * D, rot_repr, T_in + 16 * (size_t)b, fx, fy, s_T, s_sv);
        float4* o4 = reinterpret_cast<float4*>(T_new + 16 * (size_t)b);
        const float4* s4 = reinterpret_cast<const float4*>(s_T);
        o4[0] = s4[0]; o4[1] = s4[1]; o4[2] = s4[2]; o4[3] = s4[3];
    }
    __syncthreads();
    const Pose34 Tn = load_pose(s_T);
    const Pose34 G = load_pose(T_gt + 16 * (size_t)b);
    const Pose34 A = pred_is_a ? Tn : G;
    const Pose34 Bp = pred_is_a ? G : Tn;
    Pose34 P = A;
    if (mode == RAC_LOSS_DISENT_SCALED) P = load_pose(T_prev + 16 * (size_t)b);
    float acc[kNAcc];
#pragma unroll
    for (int i = 0; i < kNAcc; ++i) acc[i] = 0.f;
    const float* p = pts + 3 * (size_t)b * N;
    for (int n = tid; n < N; n += kLossThreads) {
        const float x = p[3 * n], yy = p[3 * n + 1], z = p[3 * n + 2];
        loss_point(mode, A, Bp, P, x, yy, z, acc);
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
        // d(mean over the batch)/d(per-sample loss) = 1/B (what autograd hands add_loss_bwd_kernel), then the pose update's backward
        const float w = 1.0f / (float)B;
        float gT[16];
        for (int i = 0; i < 16; ++i) gT[i] = (pred_is_a ? ga[i] : gb[i]) * w;
        pose_bwd_sample(gT, s_sv, rot_repr, grad_y + (size_t)b * D);
        // the loop's `T_CO_pred = T_CO_pred_new` (train_model.py:250): may alias T_in / T_prev — every read of the old pose by
        // this CTA happened before the barrier above, and no other CTA touches sample b
        if (T_feedback) {
            float4* f4 = reinterpret_cast<float4*>(T_feedback + 16 * (size_t)b);
            const float4* s4 = reinterpret_cast<const float4*>(s_T);
            f4[0] = s4[0]; f4[1] = s4[1]; f4[2] = s4[2]; f4[3] = s4[3];
        }
        __threadfence();
        const unsigned done = atomicAdd(counter, 1u);
        is_last = (done == (unsigned)B - 1u);
    }
    __syncthreads();
    if (is_last && warp == 0) {       // deterministic batch mean: the last CTA sums the per-sample losses in index order
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

}  // namespace rac

using namespace rac;

extern "C" int rac_pose_loss_step(const float* d_y, int rot_repr, const void* d_T_in, int t_in_is_f64, double fx, double fy,
                                  const float* d_T_gt, const float* d_T_prev, const float* d_pts, int B, int N, int loss_mode,
                                  int pred_is_a, float* d_T_new, float* d_per_sample, float* d_mean, float* d_grad_y,
                                  float* d_T_feedback, unsigned int* d_counter, rac_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RAC_REQUIRE(d_y && d_T_in && d_T_gt && d_pts && d_T_new && d_per_sample && d_grad_y && d_counter, "rac_pose_loss_step: null pointer");
    RAC_REQUIRE(rot_repr >= RAC_ROT_6D && rot_repr <= RAC_ROT_3D, "rac_pose_loss_step: unknown rot_repr %d", rot_repr);
    RAC_REQUIRE(loss_mode >= RAC_LOSS_ADD_L1 && loss_mode <= RAC_LOSS_DISENT_SCALED, "rac_pose_loss_step: unknown loss mode %d", loss_mode);
    RAC_REQUIRE(loss_mode != RAC_LOSS_DISENT_SCALED || d_T_prev, "rac_pose_loss_step: scaled mode needs T_prev");
    RAC_REQUIRE(B > 0 && N > 0, "rac_pose_loss_step: B=%d N=%d", B, N);
    RAC_REQUIRE(!d_T_feedback || !t_in_is_f64 || (const void*)d_T_feedback != d_T_in, "rac_pose_loss_step: T_feedback (f32) cannot alias an f64 T_in");
    if (t_in_is_f64) {
        RAC_MAX_CARVEOUT_ONCE(pose_loss_step_kernel<double>);
        pose_loss_step_kernel<double><<<B, kLossThreads, 0, stream>>>(d_y, rot_repr, (const double*)d_T_in, fx, fy, d_T_gt, d_T_prev, d_pts, B, N,
                                                                      loss_mode, pred_is_a, d_T_new, d_per_sample, d_mean, d_grad_y, d_T_feedback, d_counter);
    } else {
        RAC_MAX_CARVEOUT_ONCE(pose_loss_step_kernel<float>);
        pose_loss_step_kernel<float><<<B, kLossThreads, 0, stream>>>(d_y, rot_repr, (const float*)d_T_in, fx, fy, d_T_gt, d_T_prev, d_pts, B, N,
                                                                     loss_mode, pred_is_a, d_T_new, d_per_sample, d_mean, d_grad_y, d_T_feedback, d_counter);
    }
    RAC_LAUNCHED();
    return RAC_OK;
}
