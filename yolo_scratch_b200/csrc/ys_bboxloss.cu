// ys_bboxloss.cu -- stand-alone BboxLoss.forward (+ backward) on dense, already-assigned tensors.
// Replaces losses/yolo_v8_loss.py:280-310 (BboxLoss.forward), :312-324 (_df_loss) and :327-330 (bbox2dist) of the
// reference for callers that use BboxLoss directly; the fused path (ys_detection_loss) does not go through here.
//
// One launch: four lanes per anchor row (one per box side; lane 0 of the group also evaluates the CIoU term), the
// foreground test first so background rows cost one byte of traffic.  Per-block partial sums in a fixed order, the last
// block to finish reduces them in fp64 (run-to-run deterministic).
#include "ys_common.cuh"

namespace ys {

constexpr int kBlThreads = 256;

struct BboxLossArgs {
    const float *pred_dist;        // (n, 4*R)
    const float4 *pred_bboxes;     // (n) xyxy, grid units
    const float2 *anchor_points;   // (A)
    const float4 *target_bboxes;   // (n) xyxy, grid units
    const float *target_scores;    // (n, nc)
    const unsigned char *fg_mask;  // (n)
    const float *tss;              // [1] target_scores_sum
    float *loss_out;               // [2] loss_iou, loss_dfl
    float4 *grad_boxes;            // (n) d loss_iou / d pred_bboxes, or nullptr
    float *grad_dist;              // (n, 4*R) d loss_dfl / d pred_dist, or nullptr
    float2 *partials;              // (gridDim.x)
    unsigned *done;                // arrival counter, self-resetting
    int n, A, nc, R;
};

__global__ void __launch_bounds__(kBlThreads) k_bbox_loss(BboxLossArgs P) {
    __shared__ float s_sum[kBlThreads / 32][2];
    __shared__ double s_fin[kBlThreads / 32][2];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long t = (long long)blockIdx.x * kBlThreads + tid;
    const int row = (int)(t >> 2), side = (int)(t & 3);
    const int R = P.R;
    float acc_iou = 0.f, acc_dfl = 0.f;
    const bool in = row < P.n;
    const bool fg = in && P.fg_mask[row] != 0;
    const float inv_tss = 1.0f / P.tss[0];
    // R == 16 with 16-byte aligned tensors (the YOLO head): a lane's 16 bins are four 16-byte words
    const bool vec = R == 16 && ((reinterpret_cast<uintptr_t>(P.pred_dist) | reinterpret_cast<uintptr_t>(P.grad_dist)) & 15) == 0;
    if (in && !fg) {
        if (P.grad_boxes && side == 0) P.grad_boxes[row] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (P.grad_dist) {
            float *g = P.grad_dist + ((size_t)row * 4 + side) * R;
            if (vec) {
#pragma unroll
                for (int j = 0; j < 4; ++j) reinterpret_cast<float4 *>(g)[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                for (int j = 0; j < R; ++j) g[j] = 0.f;
            }
        }
    }
    // weight = target_scores.sum(-1)[fg_mask]  (:297); the four lanes of the row split the classes, fixed order
    // (all lanes take part in the shuffles; background rows carry 0)
    float w = 0.f;
    if (fg) {
        const float *ts = P.target_scores + (size_t)row * P.nc;
        for (int c = side; c < P.nc; c += 4) w += ts[c];
    }
    w += __shfl_xor_sync(0xffffffffu, w, 1);
    w += __shfl_xor_sync(0xffffffffu, w, 2);
    if (fg) {
        const float4 pb = P.pred_bboxes[row], tb = P.target_bboxes[row];
        if (side == 0) {
            float g4[4];
            const float ciou = ciou_with_grad(pb, tb, g4);                      // bbox_iou(xywh=False, CIoU=True) (:298)
            acc_iou = (1.0f - ciou) * w;                                        // ((1 - iou) * weight).sum() / tss (:299)
            if (P.grad_boxes) {
                const float k = -w * inv_tss;
                P.grad_boxes[row] = make_float4(k * g4[0], k * g4[1], k * g4[2], k * g4[3]);
            }
        }
        if (R > 1) {
            // bbox2dist(anchor_points, target_bboxes, reg_max - 1) (:303, :327-330), clamp again in _df_loss (:314)
            const float2 ap = P.anchor_points[row % P.A];
            const float raw = side == 0 ? ap.x - tb.x : side == 1 ? ap.y - tb.y : side == 2 ? tb.z - ap.x : tb.w - ap.y;
            const float hi = (float)(R - 1) - 0.01f;
            const float tt = fminf(fmaxf(raw, 0.f), hi);
            const int tl = (int)tt;
            const float wl = (float)(tl + 1) - tt, wr = 1.0f - wl;
            const float *z = P.pred_dist + ((size_t)row * 4 + side) * R;
            if (vec) {
                float zz[16];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 q = reinterpret_cast<const float4 *>(z)[j];
                    zz[4 * j] = q.x; zz[4 * j + 1] = q.y; zz[4 * j + 2] = q.z; zz[4 * j + 3] = q.w;
                }
                float m = zz[0];
#pragma unroll
                for (int j = 1; j < 16; ++j) m = fmaxf(m, zz[j]);
                float s = 0.f, ztl = 0.f, ztr = 0.f;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    zz[j] = expf(zz[j] - m);                 // same operations, same order as the scalar path below
                    s += zz[j];
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) {               // z[tl], z[tl + 1] back from exp: read them again instead
                    ztl = j == tl ? z[j] : ztl;
                    ztr = j == tl + 1 ? z[j] : ztr;
                }
                const float lse = m + logf(s);
                const float ce = (lse - ztl) * wl + (lse - ztr) * wr;
                acc_dfl = ce * 0.25f * w;
                if (P.grad_dist) {
                    float *g = P.grad_dist + ((size_t)row * 4 + side) * R;
                    const float k = 0.25f * w * inv_tss, inv_s = 1.0f / s;
#pragma unroll
                    for (int j = 0; j < 16; ++j) zz[j] = k * (zz[j] * inv_s - (j == tl ? wl : 0.f) - (j == tl + 1 ? wr : 0.f));
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        reinterpret_cast<float4 *>(g)[j] = make_float4(zz[4 * j], zz[4 * j + 1], zz[4 * j + 2], zz[4 * j + 3]);
                }
            } else {
            float m = z[0];
            for (int j = 1; j < R; ++j) m = fmaxf(m, z[j]);
            float s = 0.f;
            for (int j = 0; j < R; ++j) s += expf(z[j] - m);
            const float lse = m + logf(s);
            // F.cross_entropy(pred_dist, tl) * wl + F.cross_entropy(pred_dist, tl + 1) * wr, mean over the 4 sides (:316-324)
            const float ce = (lse - z[tl]) * wl + (lse - z[tl + 1]) * wr;
            acc_dfl = ce * 0.25f * w;
            if (P.grad_dist) {
                float *g = P.grad_dist + ((size_t)row * 4 + side) * R;
                const float k = 0.25f * w * inv_tss, inv_s = 1.0f / s;
                for (int j = 0; j < R; ++j) {
                    const float p = expf(z[j] - m) * inv_s;
                    g[j] = k * (p - (j == tl ? wl : 0.f) - (j == tl + 1 ? wr : 0.f));
                }
            }
            }
        } else if (P.grad_dist) {
            P.grad_dist[(size_t)row * 4 + side] = 0.f;
        }
    }
    acc_iou = warp_sum(acc_iou);
    acc_dfl = warp_sum(acc_dfl);
    if (lane == 0) { s_sum[warp][0] = acc_iou; s_sum[warp][1] = acc_dfl; }
    __syncthreads();
    if (tid == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int w = 0; w < kBlThreads / 32; ++w) { a += s_sum[w][0]; b += s_sum[w][1]; }
        P.partials[blockIdx.x] = make_float2(a, b);
        __threadfence();
        s_last = atomicAdd(P.done, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double a = 0.0, b = 0.0;
    for (int i = tid; i < (int)gridDim.x; i += kBlThreads) {
        const float2 v = __ldcg(P.partials + i);
        a += v.x; b += v.y;
    }
    a = warp_sum(a); b = warp_sum(b);
    if (lane == 0) { s_fin[warp][0] = a; s_fin[warp][1] = b; }
    __syncthreads();
    if (tid == 0) {
        double x = 0.0, y = 0.0;
        for (int w = 0; w < kBlThreads / 32; ++w) { x += s_fin[w][0]; y += s_fin[w][1]; }
        const double inv = 1.0 / (double)P.tss[0];
        P.loss_out[0] = (float)(x * inv);
        P.loss_out[1] = (float)(y * inv);
        *P.done = 0u;
    }
}

static inline unsigned bbox_loss_blocks(int n) { return (unsigned)(((long long)n * 4 + kBlThreads - 1) / kBlThreads); }

}  // namespace ys

using namespace ys;

extern "C" size_t ys_bbox_loss_workspace_bytes(int n) {
    return 256 + align_up((size_t)bbox_loss_blocks(n > 0 ? n : 1) * sizeof(float2), 256);
}

extern "C" int ys_bbox_loss(const float *pred_dist, const float *pred_bboxes, const float *anchor_points,
                            const float *target_bboxes, const float *target_scores, const uint8_t *fg_mask,
                            const float *target_scores_sum, int n, int A, int nc, int reg_max, float *loss_out,
                            float *grad_pred_bboxes, float *grad_pred_dist, void *workspace, size_t workspace_bytes,
                            void *stream_v) {
    YS_REQUIRE(n >= 1 && A >= 1 && nc >= 1 && reg_max >= 1, YS_ERR_BAD_ARG, "n, A, nc, reg_max must be >= 1");
    YS_REQUIRE(n % A == 0, YS_ERR_BAD_ARG, "n (%d) must be a multiple of the anchor count (%d)", n, A);
    YS_REQUIRE(pred_dist && pred_bboxes && anchor_points && target_bboxes && target_scores && fg_mask && target_scores_sum &&
                   loss_out && workspace, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE((((uintptr_t)pred_bboxes | (uintptr_t)target_bboxes | (uintptr_t)grad_pred_bboxes) & 15) == 0 &&
                   ((uintptr_t)anchor_points & 7) == 0, YS_ERR_BAD_ARG, "box pointers must be 16-byte aligned, anchor_points 8-byte");
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    YS_REQUIRE(ys_bbox_loss_workspace_bytes(n) <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu",
               ys_bbox_loss_workspace_bytes(n), workspace_bytes);
    cudaStream_t stream = (cudaStream_t)stream_v;
    BboxLossArgs P;
    P.pred_dist = pred_dist; P.pred_bboxes = (const float4 *)pred_bboxes; P.anchor_points = (const float2 *)anchor_points;
    P.target_bboxes = (const float4 *)target_bboxes; P.target_scores = target_scores; P.fg_mask = fg_mask;
    P.tss = target_scores_sum; P.loss_out = loss_out; P.grad_boxes = (float4 *)grad_pred_bboxes; P.grad_dist = grad_pred_dist;
    P.done = (unsigned *)workspace;
    P.partials = (float2 *)((char *)workspace + 256);
    P.n = n; P.A = A; P.nc = nc; P.R = reg_max;
    cudaError_t ce = cudaMemsetAsync(workspace, 0, 256, stream);
    YS_REQUIRE(ce == cudaSuccess, (int)ce, "cudaMemsetAsync: %s", cudaGetErrorString(ce));
    k_bbox_loss<<<bbox_loss_blocks(n), kBlThreads, 0, stream>>>(P);
    YS_CHECK_LAUNCH("k_bbox_loss");
    return YS_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// ys_bbox_loss_static: the same kernel for callers that cannot hand over a workspace they allocated (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The kernel's scratch -- the arrival counter and the per-block partial sums -- lives
// in a file-scope __device__ array of a fixed, modest size that every GPU loading the library holds until exit.
//
// g_bbox_loss_scratch is, in HOST code, only a stand-in for the symbol: its device address is obtained once per device,
// with that device current, through cudaGetSymbolAddress, and that address is what the kernel receives.
// All calls on one device share the array: two calls that overlap on the GPU (different streams) race on it.  The arrival
// counter must start at zero: the kernel's last block resets it, and it is cleared on the stream before every launch as
// well (a cudaMemsetAsync on the symbol's address: no allocation), so a call that died mid-way cannot poison the next.
// The partial sums need no clearing (every block writes its slot before the last block reads it).
namespace ys {
constexpr int kBlStaticMaxRows = YS_BBOX_LOSS_STATIC_MAX_ROWS;
constexpr size_t kBlStaticBytes = 256 + ((size_t)kBlStaticMaxRows * 4 / kBlThreads) * sizeof(float2);
static_assert(((size_t)kBlStaticMaxRows * 4) % kBlThreads == 0, "row limit must fill whole blocks");
static __device__ __align__(256) unsigned char g_bbox_loss_scratch[kBlStaticBytes];

// Device address of g_bbox_loss_scratch on the CURRENT device (cached per device in an atomic; a lost race repeats an
// idempotent query).
static int bbox_loss_static_scratch(void **out) {
    static std::atomic<void *> cache[256];
    int dev = 0;
    cudaError_t ce = cudaGetDevice(&dev);
    YS_REQUIRE(ce == cudaSuccess, (int)ce, "ys_bbox_loss_static: cudaGetDevice: %s", cudaGetErrorString(ce));
    void *p = (dev >= 0 && dev < 256) ? cache[dev].load(std::memory_order_acquire) : nullptr;
    if (!p) {
        ce = cudaGetSymbolAddress(&p, g_bbox_loss_scratch);
        YS_REQUIRE(ce == cudaSuccess && p != nullptr, (int)(ce ? ce : cudaErrorUnknown),
                   "ys_bbox_loss_static: cudaGetSymbolAddress: %s", cudaGetErrorString(ce));
        if (dev >= 0 && dev < 256) cache[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return YS_OK;
}
}  // namespace ys

extern "C" int ys_bbox_loss_static_max_rows(void) { return kBlStaticMaxRows; }

extern "C" int ys_bbox_loss_static(const float *pred_dist, const float *pred_bboxes, const float *anchor_points,
                                   const float *target_bboxes, const float *target_scores, const uint8_t *fg_mask,
                                   const float *target_scores_sum, int n, int A, int nc, int reg_max, float *loss_out,
                                   float *grad_pred_bboxes, float *grad_pred_dist, void *stream_v) {
    YS_REQUIRE(n >= 1 && A >= 1 && nc >= 1 && reg_max >= 1, YS_ERR_BAD_ARG, "n, A, nc, reg_max must be >= 1");
    YS_REQUIRE(n <= kBlStaticMaxRows, YS_ERR_UNSUPPORTED,
               "ys_bbox_loss_static: n = %d rows exceeds the built-in scratch (at most %d rows); use ys_bbox_loss with a "
               "caller-provided workspace", n, kBlStaticMaxRows);
    YS_REQUIRE(n % A == 0, YS_ERR_BAD_ARG, "n (%d) must be a multiple of the anchor count (%d)", n, A);
    YS_REQUIRE(pred_dist && pred_bboxes && anchor_points && target_bboxes && target_scores && fg_mask && target_scores_sum &&
                   loss_out, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE((((uintptr_t)pred_bboxes | (uintptr_t)target_bboxes | (uintptr_t)grad_pred_bboxes) & 15) == 0 &&
                   ((uintptr_t)anchor_points & 7) == 0, YS_ERR_BAD_ARG, "box pointers must be 16-byte aligned, anchor_points 8-byte");
    void *scratch = nullptr;
    const int rc = bbox_loss_static_scratch(&scratch);
    if (rc != YS_OK) return rc;
    cudaStream_t stream = (cudaStream_t)stream_v;
    BboxLossArgs P;
    P.pred_dist = pred_dist; P.pred_bboxes = (const float4 *)pred_bboxes; P.anchor_points = (const float2 *)anchor_points;
    P.target_bboxes = (const float4 *)target_bboxes; P.target_scores = target_scores; P.fg_mask = fg_mask;
    P.tss = target_scores_sum; P.loss_out = loss_out; P.grad_boxes = (float4 *)grad_pred_bboxes; P.grad_dist = grad_pred_dist;
    P.done = (unsigned *)scratch;
    P.partials = (float2 *)((char *)scratch + 256);
    P.n = n; P.A = A; P.nc = nc; P.R = reg_max;
    cudaError_t ce = cudaMemsetAsync(scratch, 0, 256, stream);
    YS_REQUIRE(ce == cudaSuccess, (int)ce, "ys_bbox_loss_static: cudaMemsetAsync: %s", cudaGetErrorString(ce));
    k_bbox_loss<<<bbox_loss_blocks(n), kBlThreads, 0, stream>>>(P);
    YS_CHECK_LAUNCH("k_bbox_loss (static scratch)");
    return YS_OK;
}
