// aux_kernels.cu — the two "next" rows of SURVEY.md §8(f) that sit either side of the hot path:
//   N3  batched pose sampling: se3_helpers.get_T_CO_init_and_gt (se3_helpers.py:76-88) for B samples in one launch;
//   N4  silhouette ring + overlay blend: visualization.create_silhouette / blend_imgs (visualization.py:19-41).
// Both are tiny streaming kernels (launch-latency bound for N3, HBM bound for N4); no tensor cores.
#include "rac_common.cuh"
#include <math.h>

namespace rac {

// ---------------------------------------------------------------------------------------------
// N3: pose composition from pre-drawn variates
// ---------------------------------------------------------------------------------------------
// The reference draws, per sample and in this order (se3_helpers.py:39-88):
//   get_random_rotation_translation: theta_gt ~ U(0,360) deg, axis = random(3) - 0.5, t_gt ~ N(0, dev_gt)^3
//   apply_small_random_rotation_translation: theta ~ U(-a,a) deg, t ~ U(-r,r)^3, axis = random(3) - 0.5
//   get_T_CW: distance = base + U(-dev, dev)
// = 12 uniforms in [0,1) and 3 standard normals.  The caller draws them (torch RNG on the device) and this kernel
// turns them into T_CO_init / T_CO_gt, in f64 like the numpy reference, written as f32 row-major 4x4.
struct PoseParams {
    double dist, dist_dev, gt_transl_dev, transl_dev, angle_dev_deg;
};

// Rodrigues (spatialmath angvec2r): identity when the angle or the axis vanishes
__device__ void angle_axis(double th, double ax, double ay, double az, double R[9]) {
    const double eps = 10.0 * 2.220446049250313e-16;
    const double n = sqrt(ax * ax + ay * ay + az * az);
    if (n < eps || fabs(th) < eps) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        return;
    }
    const double x = ax / n, y = ay / n, z = az / n;
    const double s = sin(th), c1 = 1.0 - cos(th);
    // I + s K + (1-c) K^2
    R[0] = 1.0 + c1 * (-(y * y) - z * z); R[1] = -s * z + c1 * (x * y);       R[2] = s * y + c1 * (x * z);
    R[3] = s * z + c1 * (x * y);        R[4] = 1.0 + c1 * (-(x * x) - z * z); R[5] = -s * x + c1 * (y * z);
    R[6] = -s * y + c1 * (x * z);       R[7] = s * x + c1 * (y * z);        R[8] = 1.0 + c1 * (-(x * x) - y * y);
}

__global__ void __launch_bounds__(128)
pose_sample_kernel(const float* __restrict__ u01, const float* __restrict__ n01, PoseParams P, int B,
                   float* __restrict__ T_init, float* __restrict__ T_gt) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const float* u = u01 + 12 * (size_t)b;
    const float* g = n01 + 3 * (size_t)b;
    const double kDeg = 3.141592653589793 / 180.0;
    // T_WO_gt (se3_helpers.py:59-68)
    double Rg[9];
    angle_axis(360.0 * (double)u[0] * kDeg, (double)u[1] - 0.5, (double)u[2] - 0.5, (double)u[3] - 0.5, Rg);
    const double tg[3] = {P.gt_transl_dev * (double)g[0], P.gt_transl_dev * (double)g[1], P.gt_transl_dev * (double)g[2]};
    // delta (se3_helpers.py:39-57): uniform sampling
    const double a = P.angle_dev_deg, r = P.transl_dev;
    const double th = (-a + 2.0 * a * (double)u[4]) * kDeg;
    const double td[3] = {-r + 2.0 * r * (double)u[5], -r + 2.0 * r * (double)u[6], -r + 2.0 * r * (double)u[7]};
    double Rd[9];
    angle_axis(th, (double)u[8] - 0.5, (double)u[9] - 0.5, (double)u[10] - 0.5, Rd);
    // T_WO_init = delta * T_WO_gt
    double Ri[9], ti[3];
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) Ri[3 * i + j] = Rd[3 * i] * Rg[j] + Rd[3 * i + 1] * Rg[3 + j] + Rd[3 * i + 2] * Rg[6 + j];
        ti[i] = Rd[3 * i] * tg[0] + Rd[3 * i + 1] * tg[1] + Rd[3 * i + 2] * tg[2] + td[i];
    }
    // T_CW = inv(look_at([d,0,0], 0, up z)) (se3_helpers.py:5-22,70-74): camera axes in world coordinates are
    // z = (-1,0,0), x = -(up x z)/|.| = (0,1,0), y = z x x = (0,0,-1); R_WC = [x y z] columns, so R_CW = rows x,y,z.
    const double d = P.dist + (-P.dist_dev + 2.0 * P.dist_dev * (double)u[11]);
    const double Rc[9] = {0, 1, 0, 0, 0, -1, -1, 0, 0};
    const double tc[3] = {0.0, 0.0, d};      // -R_CW * (d,0,0)
    auto emit = [&](const double* R, const double* t, float* out) {
        for (int i = 0; i < 3; ++i) {
            for (int j = 0; j < 3; ++j)
                out[4 * i + j] = (float)(Rc[3 * i] * R[j] + Rc[3 * i + 1] * R[3 + j] + Rc[3 * i + 2] * R[6 + j]);
            out[4 * i + 3] = (float)(Rc[3 * i] * t[0] + Rc[3 * i + 1] * t[1] + Rc[3 * i + 2] * t[2] + tc[i]);
        }
        out[12] = 0.f; out[13] = 0.f; out[14] = 0.f; out[15] = 1.f;
    };
    emit(Ri, ti, T_init + 16 * (size_t)b);
    emit(Rg, tg, T_gt + 16 * (size_t)b);
}

// ---------------------------------------------------------------------------------------------
// N4: silhouette ring and overlay blend
// ---------------------------------------------------------------------------------------------
// create_silhouette (visualization.py:19-31): sil = mean(img, 2) > 0; ring = clip(box5x5(sil), 0, 1) - sil (zero padded
// 'same' convolution) — the pixels within a 5x5 window of the object that are not object.  mode 0: write the ring as a
// (B,H,W) f32 mask; mode 1: copy the image and paint ring pixels with `colour`.
// One CTA = 32 x 8 output pixels; the (32+4) x (8+4) mask halo is staged in shared memory.
constexpr int kSilW = 32, kSilH = 8;

__global__ void __launch_bounds__(kSilW * kSilH)
silhouette_kernel(const float* __restrict__ img, int W, int H, int mode, float cr, float cg, float cb,
                  float* __restrict__ out) {
    __shared__ unsigned char s_m[kSilH + 4][kSilW + 4];
    const int b = blockIdx.z;
    const int x0 = blockIdx.x * kSilW, y0 = blockIdx.y * kSilH;
    const float* ib = img + 3 * (size_t)b * W * H;
    for (int i = threadIdx.x; i < (kSilH + 4) * (kSilW + 4); i += kSilW * kSilH) {
        const int ly = i / (kSilW + 4), lx = i - ly * (kSilW + 4);
        const int x = x0 + lx - 2, y = y0 + ly - 2;
        unsigned char m = 0;
        if (x >= 0 && x < W && y >= 0 && y < H) {
            const float* p = ib + 3 * ((size_t)y * W + x);
            // np.mean over the channel axis: pairwise/sequential sum of 3 values then / 3; only the sign matters
            m = ((p[0] + p[1] + p[2]) / 3.0f > 0.0f) ? 1 : 0;
        }
        s_m[ly][lx] = m;
    }
    __syncthreads();
    const int lx = threadIdx.x & (kSilW - 1), ly = threadIdx.x / kSilW;
    const int x = x0 + lx, y = y0 + ly;
    if (x >= W || y >= H) return;
    int any = 0;
#pragma unroll
    for (int dy = 0; dy < 5; ++dy)
#pragma unroll
        for (int dx = 0; dx < 5; ++dx) any |= s_m[ly + dy][lx + dx];
    const int ring = any && !s_m[ly + 2][lx + 2];
    const size_t pi = (size_t)b * W * H + (size_t)y * W + x;
    if (mode == 0) {
        out[pi] = ring ? 1.0f : 0.0f;
    } else {
        const float* p = ib + 3 * ((size_t)y * W + x);
        float* o = out + 3 * pi;
        o[0] = ring ? cr : p[0]; o[1] = ring ? cg : p[1]; o[2] = ring ? cb : p[2];
    }
}

// blend_imgs (visualization.py:34-37): both images -> uint8 (np.uint8(x * 255.0): truncation toward zero), PIL
// Image.blend = (uint8)(a + alpha * (b - a)) in f32 (truncation), / 255.0.  Output f32 (the reference returns f64 of
// the same values).  Inputs are expected in [0,1] like the reference's; quantized != 0: the inputs already hold the
// 8-bit codes (0..255 as floats) — the host wrapper quantises f64 inputs itself, in f64 like numpy does.
__global__ void __launch_bounds__(256)
blend_kernel(const float* __restrict__ a, const float* __restrict__ b, float alpha, int quantized, size_t n, float* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int qa = quantized ? (int)a[i] : (int)(unsigned char)(int)(a[i] * 255.0f);
    const int qb = quantized ? (int)b[i] : (int)(unsigned char)(int)(b[i] * 255.0f);
    const float t = (float)qa + alpha * (float)(qb - qa);
    const int q = (int)(unsigned char)(int)t;
    out[i] = (float)((double)q / 255.0);
}

}  // namespace rac

using namespace rac;

extern "C" {

int rac_sample_poses(const float* d_u01, const float* d_n01, int B, double distance, double distance_dev,
                     double gt_transl_dev, double transl_dev, double angle_dev_deg, float* d_T_init, float* d_T_gt,
                     rac_stream_t stream_) {
    RAC_REQUIRE(d_u01 && d_n01 && d_T_init && d_T_gt, "rac_sample_poses: null pointer");
    RAC_REQUIRE(B > 0, "rac_sample_poses: B=%d must be positive", B);
    PoseParams P = {distance, distance_dev, gt_transl_dev, transl_dev, angle_dev_deg};
    pose_sample_kernel<<<(B + 127) / 128, 128, 0, (cudaStream_t)stream_>>>(d_u01, d_n01, P, B, d_T_init, d_T_gt);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_silhouette(const float* d_img_hwc, int B, int W, int H, int silhouette_only, const float* h_colour3,
                   float* d_out, rac_stream_t stream_) {
    RAC_REQUIRE(d_img_hwc && d_out, "rac_silhouette: null pointer");
    RAC_REQUIRE(B > 0 && W > 0 && H > 0 && B <= 65535, "rac_silhouette: bad shape B=%d W=%d H=%d", B, W, H);
    const float cr = h_colour3 ? h_colour3[0] : 1.0f, cg = h_colour3 ? h_colour3[1] : 0.0f, cb = h_colour3 ? h_colour3[2] : 0.0f;
    dim3 grid((W + kSilW - 1) / kSilW, (H + kSilH - 1) / kSilH, B);
    RAC_REQUIRE(grid.y <= 65535, "rac_silhouette: H=%d too large", H);
    silhouette_kernel<<<grid, kSilW * kSilH, 0, (cudaStream_t)stream_>>>(d_img_hwc, W, H, silhouette_only ? 0 : 1, cr, cg, cb, d_out);
    RAC_LAUNCHED();
    return RAC_OK;
}

int rac_blend(const float* d_im1, const float* d_im2, float alpha, int quantized_inputs, size_t n, float* d_out, rac_stream_t stream_) {
    RAC_REQUIRE(d_im1 && d_im2 && d_out, "rac_blend: null pointer");
    RAC_REQUIRE(n > 0 && n < ((size_t)1 << 40), "rac_blend: bad element count");
    RAC_REQUIRE(alpha >= 0.0f && alpha <= 1.0f, "rac_blend: alpha=%f outside [0,1]", (double)alpha);
    blend_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(d_im1, d_im2, alpha, quantized_inputs, n, d_out);
    RAC_LAUNCHED();
    return RAC_OK;
}

}  // extern "C"
