// ys_api.cu -- library identity, thread-local error text, device query.
#include "ys_common.cuh"
#include <string.h>

namespace ys {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace ys

extern "C" int ys_version(void) { return 100; }
extern "C" const char *ys_last_error(void) { return ys::g_err; }
extern "C" int ys_device_info(int *sm_count, int *cc_major, int *cc_minor) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { ys::set_error("cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
    cudaDeviceProp p;
    e = cudaGetDeviceProperties(&p, dev);
    if (e != cudaSuccess) { ys::set_error("cudaGetDeviceProperties: %s", cudaGetErrorString(e)); return (int)e; }
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    return 0;
}

// ---------------------------------------------------------------------------------------------------
// Host-side test hooks: the __host__ __device__ scalar math of ys_math.cuh, callable without a GPU
// (tests/test_host_math.py).  Not part of the product surface; declared in tests only.
// ---------------------------------------------------------------------------------------------------
extern "C" float ys_host_iou_assign(const float *p, const float *q) {
    float4 a = make_float4(p[0], p[1], p[2], p[3]), b = make_float4(q[0], q[1], q[2], q[3]);
    return ys::iou_assign(a, ys::area_eps_h(a.x, a.y, a.z, a.w), b, ys::area_eps_h(b.x, b.y, b.z, b.w));
}
extern "C" float ys_host_iou_nms(const float *p, const float *q) {
    float4 a = make_float4(p[0], p[1], p[2], p[3]), b = make_float4(q[0], q[1], q[2], q[3]);
    return ys::iou_nms(a, ys::area_plain(a), b, ys::area_plain(b));
}
extern "C" int ys_host_centre_in_box(float ax, float ay, const float *q) {
    return ys::centre_in_box(ax, ay, make_float4(q[0], q[1], q[2], q[3])) ? 1 : 0;
}
extern "C" float ys_host_ciou_grad(const float *p, const float *t, float *g4) {
    return ys::ciou_with_grad(make_float4(p[0], p[1], p[2], p[3]), make_float4(t[0], t[1], t[2], t[3]), g4);
}
extern "C" void ys_host_bce(float x, float t, float *loss, float *sig) { ys::bce_logits(x, t, *loss, *sig); }
extern "C" float ys_host_align_norm(float score, float ov, float alpha, float beta, float eps) {
    return ys::align_norm(score, ov, alpha, beta, eps);
}
