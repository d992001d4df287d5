// ys_common.cuh -- host-side plumbing shared by the .cu translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <atomic>
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
//
// Anchor index space.  The per-anchor state the kernels exchange (decoded boxes, assignment, target scores) is indexed
// by a PADDED anchor index: every level's anchor count is rounded up to a multiple of 4, so that a level's first anchor,
// an image's first anchor and every 16-anchor item start on a 16-byte boundary of a 4-byte array (bulk copies need
// that) whatever h*w is.  a_off / A are in that space; a_true_off / A_true are the reference's dense numbering
// (make_anchors, losses/yolo_v8_loss.py:541-552), used only when state is exported to the caller.  Padded indices
// keep the order of the true ones, so "first maximum" tie-breaks are unchanged.  When every h*w is a multiple of 4
// the two spaces coincide.
struct LevelSet {
    int n;            // number of levels
    int A;            // anchors per image, padded space (multiple of 4)
    int A_true;       // anchors per image as the reference counts them
    int C;            // channels = 4*reg_max + nc
    int n_chunks;     // total 128-anchor chunks per image (class items of the loss kernel)
    const float *x[YS_MAX_LEVELS];
    float *g[YS_MAX_LEVELS];
    int w[YS_MAX_LEVELS];
    int hw[YS_MAX_LEVELS];
    int tma[YS_MAX_LEVELS];             // 1: rows of this level are 16-byte aligned (h*w % 4 == 0 and aligned base pointers):
                                        //    tiles come by TMA and gradients leave as float4; 0: plain-load tile loaders
    int a_off[YS_MAX_LEVELS + 1];       // first anchor of each level (padded space); a_off[n] = A
    int a_true_off[YS_MAX_LEVELS + 1];  // first anchor of each level (reference numbering); a_true_off[n] = A_true
    int chunk_off[YS_MAX_LEVELS + 1];   // first chunk of each level
    int hseg_off[YS_MAX_LEVELS + 1];    // first 16-anchor box item of each level (k_loss_stream)
    int n_hsegs;                        // total 16-anchor box items per image
    int sblk_off[YS_MAX_LEVELS + 1];    // first 256-anchor scan block of each level (k_assign, fused path)
    int n_sblk;                         // scan blocks per image
    float stride[YS_MAX_LEVELS];
};

constexpr int kChunk = 128;   // anchors per class item of the loss kernel (32 lanes x 4 anchors)

// Upper bound of the padded anchor count for a true count A (each of at most YS_MAX_LEVELS levels adds < 4).
inline int padded_anchor_bound(int A) { return (A + 3 * YS_MAX_LEVELS + 3) & ~3; }

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- per-device caches ------------------------------------------------------------------------------------------
// The library keeps no state that a result depends on.  The only process-wide data are per-device caches of facts
// that cannot change while the process lives -- "this kernel's dynamic shared-memory limit has been raised on device d"
// and the SM count of device d -- held in atomics (any thread, any device; a lost race repeats an idempotent call).
struct DeviceOnce {
    std::atomic<unsigned long long> w[4];      // bit d: done on device d (devices 0..255)
    bool test(int dev) const { return dev >= 0 && dev < 256 && ((w[dev >> 6].load(std::memory_order_acquire) >> (dev & 63)) & 1ull); }
    void set(int dev) { if (dev >= 0 && dev < 256) w[dev >> 6].fetch_or(1ull << (dev & 63), std::memory_order_release); }
};

// Dynamic shared-memory opt-in of a kernel on the CURRENT device, once per device (the attribute is sticky per device and
// function; repeating the call on every launch costs host time for nothing).
inline int optin_smem(const void *kernel, size_t bytes, DeviceOnce &once, const char *what) {
    int dev = 0;
    cudaError_t ce = cudaGetDevice(&dev);
    YS_REQUIRE(ce == cudaSuccess, (int)ce, "%s: cudaGetDevice: %s", what, cudaGetErrorString(ce));
    if (once.test(dev)) return YS_OK;
    ce = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    YS_REQUIRE(ce == cudaSuccess, (int)ce, "%s smem opt-in: %s", what, cudaGetErrorString(ce));
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    once.set(dev);
    return YS_OK;
}

// SM count of the current device (cached per device).
inline int device_sm_count(int &sms) {
    static std::atomic<int> cache[256];
    int dev = 0;
    cudaError_t ce = cudaGetDevice(&dev);
    YS_REQUIRE(ce == cudaSuccess, (int)ce, "cudaGetDevice: %s", cudaGetErrorString(ce));
    int v = (dev >= 0 && dev < 256) ? cache[dev].load(std::memory_order_relaxed) : 0;
    if (v <= 0) {
        ce = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        YS_REQUIRE(ce == cudaSuccess && v > 0, (int)(ce ? ce : cudaErrorUnknown), "cannot query the SM count: %s", cudaGetErrorString(ce));
        if (dev >= 0 && dev < 256) cache[dev].store(v, std::memory_order_relaxed);
    }
    sms = v;
    return YS_OK;
}

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
