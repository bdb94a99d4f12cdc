// rac_common.cuh — shared host/device plumbing of librac_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>
#include "../../include/rac.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "librac_b200 is written for sm_100a (B200) only"
#endif

namespace rac {

// ---- error plumbing ---------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;

#define RAC_CUDA_CHECK(expr)                                                        \
    do {                                                                            \
        cudaError_t _e = (expr);                                                    \
        if (_e != cudaSuccess) {                                                    \
            rac::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),  \
                           __FILE__, __LINE__);                                     \
            return RAC_E_CUDA;                                                      \
        }                                                                           \
    } while (0)

#define RAC_REQUIRE(cond, ...)                                                      \
    do {                                                                            \
        if (!(cond)) {                                                              \
            rac::set_error(__VA_ARGS__);                                            \
            return RAC_E_INVALID;                                                   \
        }                                                                           \
    } while (0)

// Call after every kernel launch: counts it and surfaces launch-configuration errors.
#define RAC_LAUNCHED()                                                              \
    do {                                                                            \
        rac::g_launches.fetch_add(1, std::memory_order_relaxed);                    \
        RAC_CUDA_CHECK(cudaGetLastError());                                         \
    } while (0)

// Every kernel of the library asks for the same L1 / shared-memory split (maximum shared memory, what raster_kernel
// needs): an SM does not mix carve-outs, so this is what lets the small pose / loss kernels of a captured step run on
// a parallel graph branch next to the resident raster CTAs instead of waiting for an SM to drain.
#define RAC_MAX_CARVEOUT_ONCE(kernel)                                                                       \
    do {                                                                                                    \
        static std::atomic<unsigned long long> _done{0};      /* bit per device ordinal */                  \
        int _dev = 0;                                                                                       \
        RAC_CUDA_CHECK(cudaGetDevice(&_dev));                                                               \
        const unsigned long long _bit = 1ull << (_dev & 63);                                                \
        if (!(_done.load(std::memory_order_acquire) & _bit)) {                                              \
            RAC_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,     \
                                                cudaSharedmemCarveoutMaxShared));                           \
            _done.fetch_or(_bit, std::memory_order_release);                                                \
        }                                                                                                   \
    } while (0)

// ---- device-resident mesh table ---------------------------------------------
// One primitive ("part") of a multi-material mesh (a .glb scene: renderer.py:36-42 adds every geometry of the file as
// its own pyrender mesh, each with the material the file gives it).  Shading constants are precomputed on the host.
struct PartDev {
    float base_color[4];
    float metallic, alpha2, kgeo;   // alpha2 = roughness^4, kgeo = (roughness + 1)^2 / 8 (roughness clamped to [0.04, 1])
    int has_vcolor;                 // COLOR_0 applies to this part (parts without a material, pyrender's default one)
    const uchar4* tex;              // this part's mip chain, or nullptr
    int tw, th, nmips;
    int double_sided;               // glTF doubleSided: back faces are drawn too (unlit side: the normal is not flipped)
};

struct MeshDev {
    const float*   verts;     // V*3 object-frame positions
    const int*     faces;     // T*3 vertex indices
    const float*   fnormals;  // T*3 unit face normals (object frame), 0 for degenerate faces
    const float*   uv;        // V*2 or nullptr
    const uchar4*  vcolor;    // V   or nullptr (RGBA8, COLOR_0)
    const uchar4*  tex;       // mip chain, level 0 first (RGBA8, sRGB-encoded) or nullptr
    int V, T;
    int tw, th, nmips;
    float base_color[4];
    float metallic, roughness;
    const PartDev* parts;         // multi-part meshes only (else nullptr): per-part material / texture ...
    const unsigned char* face_part;   // ... selected per face, T entries
};

constexpr int kMaxMeshes = 4096;
const MeshDev* mesh_table_dev();                  // device pointer to MeshDev[kMaxMeshes]
int mesh_host_info(rac_mesh_t h, int* V, int* T); // 0 ok

// ---- small device helpers ---------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ unsigned warp_max(unsigned v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_maxf(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace rac
