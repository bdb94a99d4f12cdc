// host_pack.cu — host-side staging transform for the gt crop the UNCHANGED loader hands to prepare_model_input.
//
// ref: image_dataloaders.py:79-83 — `real_img = np.asarray(Image.open(real_path).convert('RGB')); real_img = real_img/255.0`:
// the crop reaches prepare_model_input (train_model.py:226,243 / test_model.py:91,100) as PAGEABLE float64 numpy, 24 B per
// pixel, although it carries 3 B per pixel of information.  rac_host_pack_gt_u8 recovers the PNG codes in ONE
// multi-threaded pass over the array — q = rint(v * 255), accepted only if q / 255.0 == v exactly (the very f64 division
// the loader performed) — into a pinned staging buffer, so the crop crosses PCIe as bytes and rac_gt_norm_u8 evaluates
// normalize_img(q / 255.0) on the device, bit-identical to the reference's f64 -> f32 cast.  If any value is not an
// exact code (augmented or synthetic float crops) the call says so and the caller uploads floats instead.
// This is data staging, not a compute fallback: no normalisation, rendering, pose or loss arithmetic happens on the host.
#include "rac_common.cuh"
#include <atomic>
#include <stdlib.h>
#include <thread>
#include <vector>

namespace rac {

// One block of values: codes out, 1 if every value is an exact code.  Branch-free and table-free so that the host
// compiler vectorises it (cvttpd2dq + divpd + cmppd); the AVX2 clone is picked at run time.
#define RAC_PACK_BLOCK_BODY                                                                        \
    int ok = 1;                                                                                    \
    for (size_t i = 0; i < n; ++i) {                                                               \
        const F v = src[i];                                                                        \
        const int q = (int)(v * (F)255 + (F)0.5);                                                  \
        dst[i] = (unsigned char)q;                                                                 \
        const F r = (F)((double)q / 255.0);      /* f64: the loader's own division; f32: its cast */ \
        ok &= (int)(r == v) & (int)(v >= (F)0) & (int)(v <= (F)1);   /* NaN fails all three */    \
    }                                                                                              \
    return ok;

template <typename F>
static int pack_block_generic(const F* __restrict__ src, unsigned char* __restrict__ dst, size_t n) { RAC_PACK_BLOCK_BODY }
template <typename F>
__attribute__((target("avx2"))) static int pack_block_avx2(const F* __restrict__ src, unsigned char* __restrict__ dst, size_t n) { RAC_PACK_BLOCK_BODY }

// float64 sources on CPUs with FMA: the same acceptance test WITHOUT the division.  q / 255 is never a midpoint between two
// doubles (255 = 2^8 - 1: the quotient's binary expansion does not terminate unless q is 0 or 255), so v == RN(q / 255) holds
// exactly when |v - q / 255| < ulp(v) / 2, i.e. |255 v - q| < 127.5 ulp(v); fma(v, 255, -q) delivers 255 v - q EXACTLY (after
// the cancellation the residual has a handful of significant bits), and ulp(v) = 2^(exponent(v) - 52) comes from the bit
// pattern.  Cross-checked against the division form on every code and its neighbouring doubles (tests/test_cabi_and_host.py).
__attribute__((target("avx2,fma"))) static int pack_block_f64_fma(const double* __restrict__ src, unsigned char* __restrict__ dst, size_t n) {
    int ok = 1;
    for (size_t i = 0; i < n; ++i) {
        const double v = src[i];
        const int q = (int)(v * 255.0 + 0.5);
        dst[i] = (unsigned char)q;
        const double e = __builtin_fma(v, 255.0, -(double)q);
        unsigned long long bits;
        __builtin_memcpy(&bits, &v, 8);
        const unsigned long long ub = (bits & 0x7FF0000000000000ull) - (52ull << 52);      // bit pattern of ulp(v), v normal and >= 2^-970
        double ulp;
        __builtin_memcpy(&ulp, &ub, 8);
        const int exact = (int)(__builtin_fabs(e) < 127.5 * ulp) & (int)(v >= 0x1p-900);
        ok &= (exact | (int)(v == 0.0)) & (int)(v >= 0.0) & (int)(v <= 1.0);                // NaN fails every comparison
    }
    return ok;
}

template <typename F>
static int pack_block(const F* src, unsigned char* dst, size_t n, int isa) {
    return isa ? pack_block_avx2<F>(src, dst, n) : pack_block_generic<F>(src, dst, n);
}
template <>
int pack_block<double>(const double* src, unsigned char* dst, size_t n, int isa) {
    return isa == 2 ? pack_block_f64_fma(src, dst, n) : isa ? pack_block_avx2<double>(src, dst, n) : pack_block_generic<double>(src, dst, n);
}

template <typename F>
static void pack_range(const F* src, unsigned char* dst, size_t lo, size_t hi, int avx2, std::atomic<int>* bad) {
    constexpr size_t kBlock = 1 << 14;
    for (size_t b0 = lo; b0 < hi; b0 += kBlock) {
        if (bad->load(std::memory_order_relaxed)) return;      // another thread already found a non-code value
        const size_t nb = b0 + kBlock < hi ? kBlock : hi - b0;
        const int ok = pack_block<F>(src + b0, dst + b0, nb, avx2);
        if (!ok) { bad->store(1, std::memory_order_relaxed); return; }
    }
}

template <typename F>
static int pack_all(const F* src, size_t n, unsigned char* dst, int n_threads) {
    const int avx2 = !__builtin_cpu_supports("avx2") ? 0 : (__builtin_cpu_supports("fma") && !getenv("RAC_PACK_NO_FMA")) ? 2 : 1;
    std::atomic<int> bad{0};
    if (n_threads < 1) n_threads = 1;
    if ((size_t)n_threads > n / (1 << 16) + 1) n_threads = (int)(n / (1 << 16) + 1);
    std::vector<std::thread> pool;
    const size_t per = ((n + n_threads - 1) / n_threads + 63) & ~(size_t)63;
    for (int t = 1; t < n_threads; ++t) {
        const size_t lo = per * t, hi = lo + per < n ? lo + per : n;
        if (lo < hi) pool.emplace_back(pack_range<F>, src, dst, lo, hi, avx2, &bad);
    }
    pack_range<F>(src, dst, 0, per < n ? per : n, avx2, &bad);
    for (auto& th : pool) th.join();
    return bad.load() ? 0 : 1;
}

}  // namespace rac

extern "C" int rac_host_pack_gt_u8(const void* h_src, int src_is_f64, size_t n, unsigned char* h_dst, int n_threads, int* all_exact) {
    using namespace rac;
    RAC_REQUIRE(h_src && h_dst && all_exact, "rac_host_pack_gt_u8: null argument");
    if (n_threads <= 0) {
        n_threads = (int)std::thread::hardware_concurrency();
        if (n_threads > 16) n_threads = 16;        // memory-bound: more threads than memory channels buy nothing
    }
    *all_exact = src_is_f64 ? pack_all<double>((const double*)h_src, n, h_dst, n_threads)
                            : pack_all<float>((const float*)h_src, n, h_dst, n_threads);
    return RAC_OK;
}
