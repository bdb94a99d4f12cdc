// pose_update_dev.cuh — per-sample device code of steps 3+4 (rotation heads, SE(3) update), shared by the stand-alone
// kernels of pose_update.cu and the fused post-CNN kernel of pose_loss_step.cu.  See pose_update.cu for the references.
#pragma once
#include "rac_common.cuh"
#include <math.h>

namespace rac {

// saved-for-backward layout (RAC_POSE_SAVED_FLOATS = 48 floats per sample)
//  [0..8]   R_k (f32)            [9] z_k  [10] ax = vx/fx + x_k/z_k  [11] ay = vy/fy + x_k/z_k
//  [12] 1/fx [13] 1/fy [14] z_kp1
//  [15..]   head specific: 6D: a(3) b(3);  SVD: U'(9) V'(9) p(3);  3D: v(3) eps
constexpr int kSavedHead = 15;

struct Mat3 { double m[9]; };

__device__ __forceinline__ void cross3(const double* a, const double* b, double* c) {
    c[0] = a[1] * b[2] - a[2] * b[1];
    c[1] = a[2] * b[0] - a[0] * b[2];
    c[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// ---- 6D (ref: rotation_representation.py:4-19), columns of R = (x, y, z) -----------------
__device__ __forceinline__ void rot6d_fwd(const double* a, const double* b, double* R) {
    const double na = sqrt(dot3(a, a));
    double x[3] = {a[0] / na, a[1] / na, a[2] / na};   // no epsilon, like the reference (Q9)
    double z[3];
    cross3(x, b, z);
    const double nz = sqrt(dot3(z, z));
    z[0] /= nz; z[1] /= nz; z[2] /= nz;
    double y[3];
    cross3(z, x, y);
    for (int i = 0; i < 3; ++i) { R[3 * i] = x[i]; R[3 * i + 1] = y[i]; R[3 * i + 2] = z[i]; }
}

__device__ __forceinline__ void rot6d_bwd(const double* a, const double* b, const double* G, double* ga, double* gb) {
    const double na = sqrt(dot3(a, a));
    double x[3] = {a[0] / na, a[1] / na, a[2] / na};
    double zr[3];
    cross3(x, b, zr);
    const double nz = sqrt(dot3(zr, zr));
    double z[3] = {zr[0] / nz, zr[1] / nz, zr[2] / nz};
    double gx[3], gy[3], gz[3];
    for (int i = 0; i < 3; ++i) { gx[i] = G[3 * i]; gy[i] = G[3 * i + 1]; gz[i] = G[3 * i + 2]; }
    // y = z x x  ->  gz += x x gy ; gx += gy x z
    double t[3];
    cross3(x, gy, t);  gz[0] += t[0]; gz[1] += t[1]; gz[2] += t[2];
    cross3(gy, z, t);  gx[0] += t[0]; gx[1] += t[1]; gx[2] += t[2];
    // z = zr/|zr|
    const double zg = dot3(z, gz);
    double gzr[3] = {(gz[0] - z[0] * zg) / nz, (gz[1] - z[1] * zg) / nz, (gz[2] - z[2] * zg) / nz};
    // zr = x x b  ->  gx += b x gzr ; gb = gzr x x
    cross3(b, gzr, t); gx[0] += t[0]; gx[1] += t[1]; gx[2] += t[2];
    cross3(gzr, x, gb);
    // x = a/|a|
    const double xg = dot3(x, gx);
    for (int i = 0; i < 3; ++i) ga[i] = (gx[i] - x[i] * xg) / na;
}

// ---- SVD (ref: rotation_representation.py:22-39) ---------------------------------------------
// R = U diag(1,1,det(U V^T)) V^T only needs the two leading singular pairs:
//   R = u1 v1^T + u2 v2^T + (u1 x u2)(v1 x v2)^T.
// V, sigma^2 come from a cyclic Jacobi eigen-solve of M^T M.  Outputs U' = [u1 u2 u1xu2],
// V' = [v1 v2 v1xv2] (both det +1, columns) and p = diag(U'^T M V') (p3 carries the sign).
__device__ __forceinline__ void jacobi_rotate(double* A, double* V, int p, int q) {
    const double apq = A[3 * p + q];
    if (fabs(apq) < 1e-300) return;
    const double theta = (A[3 * q + q] - A[3 * p + p]) / (2.0 * apq);
    const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
    const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
    for (int k = 0; k < 3; ++k) {   // A <- A J
        const double akp = A[3 * k + p], akq = A[3 * k + q];
        A[3 * k + p] = c * akp - s * akq;
        A[3 * k + q] = s * akp + c * akq;
    }
    for (int k = 0; k < 3; ++k) {   // A <- J^T A
        const double apk = A[3 * p + k], aqk = A[3 * q + k];
        A[3 * p + k] = c * apk - s * aqk;
        A[3 * q + k] = s * apk + c * aqk;
    }
    for (int k = 0; k < 3; ++k) {   // V <- V J
        const double vkp = V[3 * k + p], vkq = V[3 * k + q];
        V[3 * k + p] = c * vkp - s * vkq;
        V[3 * k + q] = s * vkp + c * vkq;
    }
}

__device__ __forceinline__ void svd_orth_fwd(const double* M, double* R, double* Up, double* Vp, double* p) {
    double A[9], V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) A[3 * i + j] = M[i] * M[j] + M[3 + i] * M[3 + j] + M[6 + i] * M[6 + j];
    for (int sweep = 0; sweep < 12; ++sweep) {
        const double off = A[1] * A[1] + A[2] * A[2] + A[5] * A[5];
        const double dia = A[0] * A[0] + A[4] * A[4] + A[8] * A[8];
        if (off <= 1e-34 * dia) break;
        jacobi_rotate(A, V, 0, 1);
        jacobi_rotate(A, V, 0, 2);
        jacobi_rotate(A, V, 1, 2);
    }
    // order the eigenpairs by decreasing eigenvalue; only the top two are used
    int i0 = 0, i1 = 1, i2 = 2;
    double l0 = A[0], l1 = A[4], l2 = A[8];
    if (l0 < l1) { double t = l0; l0 = l1; l1 = t; int k = i0; i0 = i1; i1 = k; }
    if (l0 < l2) { double t = l0; l0 = l2; l2 = t; int k = i0; i0 = i2; i2 = k; }
    if (l1 < l2) { double t = l1; l1 = l2; l2 = t; int k = i1; i1 = i2; i2 = k; }
    double v1[3] = {V[i0], V[3 + i0], V[6 + i0]}, v2[3] = {V[i1], V[3 + i1], V[6 + i1]}, v3[3];
    cross3(v1, v2, v3);
    double u1[3], u2[3], u3[3];
    for (int i = 0; i < 3; ++i) {
        u1[i] = M[3 * i] * v1[0] + M[3 * i + 1] * v1[1] + M[3 * i + 2] * v1[2];
        u2[i] = M[3 * i] * v2[0] + M[3 * i + 1] * v2[1] + M[3 * i + 2] * v2[2];
    }
    double n1 = sqrt(dot3(u1, u1));
    if (n1 > 0.0) { u1[0] /= n1; u1[1] /= n1; u1[2] /= n1; } else { u1[0] = 1.0; u1[1] = 0.0; u1[2] = 0.0; }
    const double pr = dot3(u1, u2);
    u2[0] -= pr * u1[0]; u2[1] -= pr * u1[1]; u2[2] -= pr * u1[2];
    double n2 = sqrt(dot3(u2, u2));
    if (n2 > 1e-150) { u2[0] /= n2; u2[1] /= n2; u2[2] /= n2; }
    else {
        // rank <= 1: any unit vector orthogonal to u1 (the reference's LAPACK choice is arbitrary too)
        double e[3] = {0, 0, 0};
        e[fabs(u1[0]) < 0.6 ? 0 : 1] = 1.0;
        cross3(u1, e, u2);
        n2 = sqrt(dot3(u2, u2));
        u2[0] /= n2; u2[1] /= n2; u2[2] /= n2;
        n2 = 0.0;
    }
    cross3(u1, u2, u3);
    for (int i = 0; i < 3; ++i) {
        Up[3 * i] = u1[i]; Up[3 * i + 1] = u2[i]; Up[3 * i + 2] = u3[i];
        Vp[3 * i] = v1[i]; Vp[3 * i + 1] = v2[i]; Vp[3 * i + 2] = v3[i];
    }
    // p_j = u_j^T M v_j
    double Mv3[3];
    for (int i = 0; i < 3; ++i) Mv3[i] = M[3 * i] * v3[0] + M[3 * i + 1] * v3[1] + M[3 * i + 2] * v3[2];
    p[0] = n1; p[1] = n2; p[2] = dot3(u3, Mv3);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[3 * i + j] = u1[i] * v1[j] + u2[i] * v2[j] + u3[i] * v3[j];
}

// dL/dM = U' Z V'^T with Z_ij = (B_ij - B_ji)/(p_i + p_j), B = U'^T G V'   
__device__ __forceinline__ void svd_orth_bwd(const double* Up, const double* Vp, const double* p, const double* G, double* gM) {
    double Bm[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k)
                for (int l = 0; l < 3; ++l) s += Up[3 * k + i] * G[3 * k + l] * Vp[3 * l + j];
            Bm[3 * i + j] = s;
        }
    double Z[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const double den = p[i] + p[j];
            Z[3 * i + j] = (i == j) ? 0.0 : (Bm[3 * i + j] - Bm[3 * j + i]) / den;
        }
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k)
                for (int l = 0; l < 3; ++l) s += Up[3 * i + k] * Z[3 * k + l] * Vp[3 * j + l];
            gM[3 * i + j] = s;
        }
}

// ---- axis-angle exp map (ref: rotation_representation.py:41-118) ------------------------------
__device__ __forceinline__ void expmap_fwd(const double* v, double eps, double* R) {
    const double nr = dot3(v, v);
    const double th = sqrt(nr < eps ? eps : nr);   // Q8: theta^2 is clamped at eps
    const double inv = 1.0 / th;
    const double f1 = inv * sin(th), f2 = inv * inv * (1.0 - cos(th));
    const double x = v[0], y = v[1], z = v[2];
    const double Km[9] = {0, -z, y, z, 0, -x, -y, x, 0};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double k2 = 0.0;
            for (int k = 0; k < 3; ++k) k2 += Km[3 * i + k] * Km[3 * k + j];
            R[3 * i + j] = f1 * Km[3 * i + j] + f2 * k2 + (i == j ? 1.0 : 0.0);
        }
}

__device__ __forceinline__ void expmap_bwd(const double* v, double eps, const double* G, double* gv) {
    const double nr = dot3(v, v);
    const bool pass = nr >= eps;   // torch.clamp(min) passes the gradient where x >= min
    const double th = sqrt(pass ? nr : eps);
    const double inv = 1.0 / th;
    const double sn = sin(th), cs = cos(th);
    const double f1 = inv * sn, f2 = inv * inv * (1.0 - cs);
    const double x = v[0], y = v[1], z = v[2];
    const double Km[9] = {0, -z, y, z, 0, -x, -y, x, 0};
    double K2[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s += Km[3 * i + k] * Km[3 * k + j];
            K2[3 * i + j] = s;
        }
    double gf1 = 0.0, gf2 = 0.0;
    for (int i = 0; i < 9; ++i) { gf1 += G[i] * Km[i]; gf2 += G[i] * K2[i]; }
    // dL/dK = f1 G + f2 (G K^T + K^T G)
    double gK[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s += G[3 * i + k] * Km[3 * j + k] + Km[3 * k + i] * G[3 * k + j];
            gK[3 * i + j] = f1 * G[3 * i + j] + f2 * s;
        }
    gv[0] = gK[7] - gK[5];
    gv[1] = gK[2] - gK[6];
    gv[2] = gK[3] - gK[1];
    if (pass) {
        const double df1 = (th * cs - sn) * inv * inv;
        const double df2 = (th * sn - 2.0 * (1.0 - cs)) * inv * inv * inv;
        const double gth = gf1 * df1 + gf2 * df2;
        gv[0] += gth * x * inv; gv[1] += gth * y * inv; gv[2] += gth * z * inv;
    }
}

__device__ __forceinline__ int head_dim(int rot_repr) { return rot_repr == RAC_ROT_SVD ? 9 : (rot_repr == RAC_ROT_6D ? 6 : 3); }

__device__ __forceinline__ void head_fwd(int rot_repr, const float* __restrict__ in, double eps, double* R, float* __restrict__ saved_head) {
    if (rot_repr == RAC_ROT_6D) {
        double a[3] = {in[0], in[1], in[2]}, b[3] = {in[3], in[4], in[5]};
        rot6d_fwd(a, b, R);
        if (saved_head) for (int i = 0; i < 6; ++i) saved_head[i] = in[i];
    } else if (rot_repr == RAC_ROT_SVD) {
        double M[9], Up[9], Vp[9], p[3];
        for (int i = 0; i < 9; ++i) M[i] = in[i];
        svd_orth_fwd(M, R, Up, Vp, p);
        if (saved_head) {
            for (int i = 0; i < 9; ++i) { saved_head[i] = (float)Up[i]; saved_head[9 + i] = (float)Vp[i]; }
            for (int i = 0; i < 3; ++i) saved_head[18 + i] = (float)p[i];
        }
    } else {
        double v[3] = {in[0], in[1], in[2]};
        expmap_fwd(v, eps, R);
        if (saved_head) { saved_head[0] = in[0]; saved_head[1] = in[1]; saved_head[2] = in[2]; saved_head[3] = (float)eps; }
    }
}

__device__ __forceinline__ void head_bwd(int rot_repr, const float* __restrict__ saved_head, const double* G, float* __restrict__ gin) {
    if (rot_repr == RAC_ROT_6D) {
        double a[3] = {saved_head[0], saved_head[1], saved_head[2]}, b[3] = {saved_head[3], saved_head[4], saved_head[5]};
        double ga[3], gb[3];
        rot6d_bwd(a, b, G, ga, gb);
        for (int i = 0; i < 3; ++i) { gin[i] = (float)ga[i]; gin[3 + i] = (float)gb[i]; }
    } else if (rot_repr == RAC_ROT_SVD) {
        double Up[9], Vp[9], p[3], gM[9];
        for (int i = 0; i < 9; ++i) { Up[i] = saved_head[i]; Vp[i] = saved_head[9 + i]; }
        for (int i = 0; i < 3; ++i) p[i] = saved_head[18 + i];
        svd_orth_bwd(Up, Vp, p, G, gM);
        for (int i = 0; i < 9; ++i) gin[i] = (float)gM[i];
    } else {
        double v[3] = {saved_head[0], saved_head[1], saved_head[2]}, gv[3];
        expmap_bwd(v, (double)saved_head[3], G, gv);
        for (int i = 0; i < 3; ++i) gin[i] = (float)gv[i];
    }
}

// One sample of calculate_T_CO_pred (rotation_representation.py:131-184): yb = the head output (hd + 3 floats), Ti = the
// 4x4 input pose (f32 or f64, Q6).  out: the 16 floats of T_new; sv: RAC_POSE_SAVED_FLOATS floats for the backward.
template <typename TIN>
__device__ __forceinline__ void pose_fwd_sample(const float* __restrict__ yb, int rot_repr, const TIN* __restrict__ Ti, double fx, double fy,
                                                float* out, float* sv) {
    const int hd = head_dim(rot_repr);
    double dR[9];
    head_fwd(rot_repr, yb, 1e-4, dR, sv + kSavedHead);
    // ref: rotation_representation.py:175 — R_k.float()
    float Rk[9] = {(float)Ti[0], (float)Ti[1], (float)Ti[2], (float)Ti[4], (float)Ti[5], (float)Ti[6], (float)Ti[8], (float)Ti[9], (float)Ti[10]};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            out[4 * i + j] = (float)(dR[3 * i] * (double)Rk[j] + dR[3 * i + 1] * (double)Rk[3 + j] + dR[3 * i + 2] * (double)Rk[6 + j]);
    const float vx = yb[hd], vy = yb[hd + 1], vz = yb[hd + 2];
    // ref: rotation_representation.py:169-174; Q1: y_k reads the x translation; Q6: fp64 when T_in is fp64
    float z_kp1, x_kp1, y_kp1, zk_f, ax_f, ay_f;
    if (sizeof(TIN) == 8) {
        const double z_k = (double)Ti[11], x_k = (double)Ti[3], y_k = (double)Ti[3];
        const double zn = (double)vz * z_k;
        // vx/fx is f32 tensor / 0-dim f64 tensor = f32 in torch; the sum with the f64 ratio promotes
        const double ax = (double)__fdiv_rn(vx, (float)fx) + x_k / z_k;
        const double ay = (double)__fdiv_rn(vy, (float)fy) + y_k / z_k;
        z_kp1 = (float)zn; x_kp1 = (float)(ax * zn); y_kp1 = (float)(ay * zn);
        zk_f = (float)z_k; ax_f = (float)ax; ay_f = (float)ay;
    } else {
        const float z_k = (float)Ti[11], x_k = (float)Ti[3], y_k = (float)Ti[3];
        const float ffx = (float)fx, ffy = (float)fy;
        const float zn = __fmul_rn(vz, z_k);
        const float ax = __fadd_rn(__fdiv_rn(vx, ffx), __fdiv_rn(x_k, z_k));
        const float ay = __fadd_rn(__fdiv_rn(vy, ffy), __fdiv_rn(y_k, z_k));
        z_kp1 = zn; x_kp1 = __fmul_rn(ax, zn); y_kp1 = __fmul_rn(ay, zn);
        zk_f = z_k; ax_f = ax; ay_f = ay;
    }
    out[3] = x_kp1; out[7] = y_kp1; out[11] = z_kp1;
    out[12] = 0.f; out[13] = 0.f; out[14] = 0.f; out[15] = 1.f;   // Q3
    for (int i = 0; i < 9; ++i) sv[i] = Rk[i];
    sv[9] = zk_f; sv[10] = ax_f; sv[11] = ay_f;
    sv[12] = (float)(1.0 / fx); sv[13] = (float)(1.0 / fy); sv[14] = z_kp1;
}

// Backward of one sample: g = d(loss)/d(T_new) (16 floats), sv = what pose_fwd_sample saved -> gyb (hd + 3 floats).
__device__ __forceinline__ void pose_bwd_sample(const float* __restrict__ g, const float* __restrict__ sv, int rot_repr, float* __restrict__ gyb) {
    const int hd = head_dim(rot_repr);
    // R' = dR R_k  ->  dL/ddR = G_R R_k^T
    double GdR[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            GdR[3 * i + j] = (double)g[4 * i] * sv[3 * j] + (double)g[4 * i + 1] * sv[3 * j + 1] + (double)g[4 * i + 2] * sv[3 * j + 2];
    head_bwd(rot_repr, sv + kSavedHead, GdR, gyb);
    const double gxp = g[3], gyp = g[7], gzp = g[11];
    const double z_k = sv[9], ax = sv[10], ay = sv[11], ifx = sv[12], ify = sv[13], z_kp1 = sv[14];
    gyb[hd] = (float)(gxp * z_kp1 * ifx);
    gyb[hd + 1] = (float)(gyp * z_kp1 * ify);
    gyb[hd + 2] = (float)((gzp + gxp * ax + gyp * ay) * z_k);
}

}  // namespace rac
