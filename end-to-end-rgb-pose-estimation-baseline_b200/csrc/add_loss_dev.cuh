// add_loss_dev.cuh — per-point / per-sample device code of step 5 (point-matching losses), shared by add_loss.cu and the
// fused post-CNN kernel of pose_loss_step.cu.  See add_loss.cu for the references.
#pragma once
#include "rac_common.cuh"

namespace rac {

constexpr int kLossThreads = 128;
constexpr int kNAcc = 17;

struct Pose34 {
    float r[9];
    float t[3];
};

__device__ __forceinline__ Pose34 load_pose(const float* __restrict__ T) {
    Pose34 p;
    p.r[0] = T[0]; p.r[1] = T[1]; p.r[2] = T[2];  p.t[0] = T[3];
    p.r[3] = T[4]; p.r[4] = T[5]; p.r[5] = T[6];  p.t[1] = T[7];
    p.r[6] = T[8]; p.r[7] = T[9]; p.r[8] = T[10]; p.t[2] = T[11];
    return p;
}

// R p + t, the reference's `T[..., :3, :3] @ pts + T[..., :3, [-1]]` (loss.py:77)
__device__ __forceinline__ void xform(const float* r, const float* t, float x, float y, float z, float* o) {
    o[0] = fmaf(r[2], z, fmaf(r[1], y, r[0] * x)) + t[0];
    o[1] = fmaf(r[5], z, fmaf(r[4], y, r[3] * x)) + t[1];
    o[2] = fmaf(r[8], z, fmaf(r[7], y, r[6] * x)) + t[2];
}
__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

// accumulators: [0] loss numerator, [1..9] S_ij = sum sign(d_i) p_j (rotation part),
// [10..12] s_i = sum sign(d_i) (translation part), [13] scaled-mode denominator numerator, [14..16] spare
__device__ __forceinline__ void loss_point(int mode, const Pose34& A, const Pose34& Bp, const Pose34& P, float x, float y, float z, float* acc) {
        if (mode == RAC_LOSS_ADD_L1 || mode == RAC_LOSS_ADD_L2) {
            // a = TCO_gt, b = TCO_pred: d = a(p) - b(p)
            float pa[3], pb[3];
            xform(A.r, A.t, x, y, z, pa);
            xform(Bp.r, Bp.t, x, y, z, pb);
            const float d0 = pa[0] - pb[0], d1 = pa[1] - pb[1], d2 = pa[2] - pb[2];
            float w0, w1, w2;
            if (mode == RAC_LOSS_ADD_L1) {
                acc[0] += fabsf(d0) + fabsf(d1) + fabsf(d2);
                w0 = sgn(d0); w1 = sgn(d1); w2 = sgn(d2);
            } else {
                const float dist = sqrtf(d0 * d0 + d1 * d1 + d2 * d2);
                acc[0] += dist;
                const float inv = dist > 0.f ? 1.0f / dist : 0.f;
                w0 = d0 * inv; w1 = d1 * inv; w2 = d2 * inv;
            }
            acc[1] += w0 * x; acc[2] += w0 * y; acc[3] += w0 * z;
            acc[4] += w1 * x; acc[5] += w1 * y; acc[6] += w1 * z;
            acc[7] += w2 * x; acc[8] += w2 * y; acc[9] += w2 * z;
            acc[10] += w0; acc[11] += w1; acc[12] += w2;
        } else {
            // a = "T_CO_pred" parameter, b = "T_CO_gt" parameter (loss.py:81): pieces of a pasted into clones of b
            float pb[3], q[3];
            xform(Bp.r, Bp.t, x, y, z, pb);
            // rotation term: [R_a | t_b]
            xform(A.r, Bp.t, x, y, z, q);
            const float r0 = pb[0] - q[0], r1 = pb[1] - q[1], r2 = pb[2] - q[2];
            acc[0] += fabsf(r0) + fabsf(r1) + fabsf(r2);
            const float w0 = sgn(r0), w1 = sgn(r1), w2 = sgn(r2);
            acc[1] += w0 * x; acc[2] += w0 * y; acc[3] += w0 * z;
            acc[4] += w1 * x; acc[5] += w1 * y; acc[6] += w1 * z;
            acc[7] += w2 * x; acc[8] += w2 * y; acc[9] += w2 * z;
            // xy-translation term: [R_b | (t_a.x, t_a.y, t_b.z)]
            const float tt[3] = {A.t[0], A.t[1], Bp.t[2]};
            xform(Bp.r, tt, x, y, z, q);
            const float t0 = pb[0] - q[0], t1 = pb[1] - q[1];
            acc[0] += fabsf(t0) + fabsf(t1) + fabsf(pb[2] - q[2]);
            acc[10] += sgn(t0); acc[11] += sgn(t1);
            // depth term: [R_b | (t_b.x, t_b.y, t_a.z)]
            const float td[3] = {Bp.t[0], Bp.t[1], A.t[2]};
            xform(Bp.r, td, x, y, z, q);
            const float e2 = pb[2] - q[2];
            acc[0] += fabsf(pb[0] - q[0]) + fabsf(pb[1] - q[1]) + fabsf(e2);
            acc[12] += sgn(e2);
            if (mode == RAC_LOSS_DISENT_SCALED) {
                // scaling = ADD_L1(T_gt, T_prev) (loss.py:116-117), no gradient
                xform(P.r, P.t, x, y, z, q);
                acc[13] += fabsf(pb[0] - q[0]) + fabsf(pb[1] - q[1]) + fabsf(pb[2] - q[2]);
            }
        }
}

// Totals of one sample (tot[0..13], summed over its N points) -> per-sample loss and d(loss)/dT_a, d(loss)/dT_b (16 floats each).
__device__ __forceinline__ void loss_finish(int mode, const float* tot, int N, float& loss_out, float* ga, float* gb) {
        const float inv3n = 1.0f / (3.0f * (float)N), invn = 1.0f / (float)N;
        float loss, gscale;
        if (mode == RAC_LOSS_ADD_L2) { loss = tot[0] * invn; gscale = invn; }
        else { loss = tot[0] * inv3n; gscale = inv3n; }
        if (mode == RAC_LOSS_DISENT_SCALED) {
            const float scaling = fmaxf(tot[13] * inv3n, 0.01f);   // loss.py:119
            loss = loss / scaling;
            gscale = gscale / scaling;
        }
        loss_out = loss;
        {
            for (int i = 0; i < 16; ++i) { ga[i] = 0.f; gb[i] = 0.f; }
            if (mode == RAC_LOSS_ADD_L1 || mode == RAC_LOSS_ADD_L2) {
                for (int i = 0; i < 3; ++i) {
                    for (int j = 0; j < 3; ++j) { ga[4 * i + j] = tot[1 + 3 * i + j] * gscale; gb[4 * i + j] = -ga[4 * i + j]; }
                    ga[4 * i + 3] = tot[10 + i] * gscale; gb[4 * i + 3] = -ga[4 * i + 3];
                }
            } else {
                // d = b(p) - [pieces of a](p): rotation of a gets -S, rotation of b gets +S;
                // translations: b gets +s, a gets -s (the R_b terms of the xy/z parts cancel exactly)
                for (int i = 0; i < 3; ++i) {
                    for (int j = 0; j < 3; ++j) { gb[4 * i + j] = tot[1 + 3 * i + j] * gscale; ga[4 * i + j] = -gb[4 * i + j]; }
                    gb[4 * i + 3] = tot[10 + i] * gscale; ga[4 * i + 3] = -gb[4 * i + 3];
                }
            }
        }
}

}  // namespace rac
