// ys_common.cuh -- host-side plumbing shared by the .cu translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/ys_b200.h"
#include "ys_math.cuh"

namespace ys {

void set_error(const char *fmt, ...);   // defined in ys_api.cu (thread-local buffer)

#define YS_REQUIRE(cond, code, ...)            \
    do {                                       \
        if (!(cond)) {                         \
            ::ys::set_error(__VA_ARGS__);      \
            return (code);                     \
        }                                      \
    } while (0)

#define YS_CHECK_LAUNCH(what)                                                        \
    do {                                                                             \
        cudaError_t e__ = cudaGetLastError();                                        \
        if (e__ != cudaSuccess) {                                                    \
            ::ys::set_error("%s: %s", (what), cudaGetErrorString(e__));              \
            return (int)e__;                                                         \
        }                                                                            \
    } while (0)

// The head levels, passed by value as a kernel parameter.
struct LevelSet {
    int n;            // number of levels
    int A;            // total anchors per image
    int C;            // channels = 4*reg_max + nc
    int n_chunks;     // total 128-anchor chunks per image (loss kernel work items)
    const float *x[YS_MAX_LEVELS];
    float *g[YS_MAX_LEVELS];
    int w[YS_MAX_LEVELS];
    int hw[YS_MAX_LEVELS];
    int a_off[YS_MAX_LEVELS + 1];       // first anchor of each level; a_off[n] = A
    int chunk_off[YS_MAX_LEVELS + 1];   // first chunk of each level
    int seg_off[YS_MAX_LEVELS + 1];     // first 32-anchor segment of each level (box role of the loss kernel)
    int n_segs;                         // total 32-anchor segments per image
    int hseg_off[YS_MAX_LEVELS + 1];    // first 16-anchor box item of each level (k_loss_stream)
    int n_hsegs;                        // total 16-anchor box items per image
    int sblk_off[YS_MAX_LEVELS + 1];    // first 256-anchor scan block of each level (k_assign, fused path)
    int n_sblk;                         // scan blocks per image
    float stride[YS_MAX_LEVELS];
};

constexpr int kChunk = 128;   // anchors per class-role work item of the loss kernel (32 lanes x float4)
constexpr int kSeg = 32;      // anchors per box-role work item (one lane per anchor, one 128-byte line per channel row)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over the caller's workspace.
struct Carver {
    char *base;
    size_t off, cap;
    Carver(void *p, size_t bytes) : base((char *)p), off(0), cap(bytes) {}
    template <typename T> T *take(size_t n) {
        off = align_up(off, 256);
        T *r = (T *)(base + off);
        off += n * sizeof(T);
        return r;
    }
    bool ok() const { return off <= cap; }
};

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
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ unsigned long long warp_max(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
        v = t > v ? t : v;
    }
    return v;
}

}  // namespace ys
