// ys_decode.cu -- inference decode (DFL + dist2bbox + sigmoid) and the IoU utility entry points.
// Replaces models/yolo11_modules.py:180-192, :248-266, :303-312, utils/metrics.py:49-81,
// train_yolo11_cuda.py:402-437 and the forward of losses/yolo_v8_loss.py:12-61; see include/ys_b200.h.
#include <string.h>
#include "ys_assign.cuh"

namespace ys {

constexpr int kDecThreads = 128;

// Channels-first output (B, 4+nc, A), what Detect.inference returns (models/yolo11_modules.py:263-266).
// One block = 128 consecutive anchors of one level of one image; thread = 4 anchors; warp w = box side w and a
// quarter of the classes (same decomposition as the loss kernel); per-channel float4 stores.
__global__ void __launch_bounds__(kDecThreads) k_decode_infer(LevelSet L, int nc, float *__restrict__ out) {
    __shared__ float s_d[4][kChunk];
    const int b = blockIdx.y, chunk = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && chunk >= L.chunk_off[k]) l = k;
    const int hw = L.hw[l], w = L.w[l];
    const float st = L.stride[l];
    const int c0 = (chunk - L.chunk_off[l]) * kChunk;
    const int i0 = c0 + lane * 4;
    const bool live = i0 < hw;
    const float *xin = L.x[l] + (size_t)b * L.C * hw + i0;
    const int a0 = L.a_off[l] + i0;
    const int R = 4 + nc;
    const int A = L.A;

    float dm[4] = {0.f, 0.f, 0.f, 0.f};
    if (live) {
        float4 z[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) z[j] = __ldg(reinterpret_cast<const float4 *>(xin + (size_t)(warp * 16 + j) * hw));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float zz[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) zz[j] = (&z[j].x)[k];
            float m2, inv_s;
            dm[k] = dfl_stats16(zz, m2, inv_s);
        }
    }
    *reinterpret_cast<float4 *>(&s_d[warp][lane * 4]) = make_float4(dm[0], dm[1], dm[2], dm[3]);
    __syncthreads();
    if (live && warp == 0) {
        float o[4][4];   // [cx,cy,w,h][k]
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = i0 + k;
            const int iy = i / w, ix = i - iy * w;
            const float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
            const float x1 = gx - s_d[0][lane * 4 + k], y1 = gy - s_d[1][lane * 4 + k];
            const float x2 = gx + s_d[2][lane * 4 + k], y2 = gy + s_d[3][lane * 4 + k];
            o[0][k] = (x1 + x2) / 2 * st;      // dist2bbox(xywh=True) * strides  (models/yolo11_modules.py:263, :309-311)
            o[1][k] = (y1 + y2) / 2 * st;
            o[2][k] = (x2 - x1) * st;
            o[3][k] = (y2 - y1) * st;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c)
            *reinterpret_cast<float4 *>(out + ((size_t)b * R + c) * A + a0) = make_float4(o[c][0], o[c][1], o[c][2], o[c][3]);
    }
    if (live) {
        const int per = (nc + 3) >> 2;
        const int c_lo = warp * per, c_hi = min(nc, c_lo + per);
        const float *xc = xin + (size_t)64 * hw;
#pragma unroll 4
        for (int c = c_lo; c < c_hi; ++c) {
            const float4 xv = __ldg(reinterpret_cast<const float4 *>(xc + (size_t)c * hw));
            const float4 s = make_float4(sigmoid_fast(xv.x), sigmoid_fast(xv.y), sigmoid_fast(xv.z), sigmoid_fast(xv.w));
            *reinterpret_cast<float4 *>(out + ((size_t)b * R + 4 + c) * A + a0) = s;
        }
    }
}

// The same decode for head shapes the vector kernel cannot take (a level with h*w % 4 != 0, or pointers that are not
// 16-byte aligned -- the reference accepts any image size, models/yolo11_modules.py:248-266): thread = one anchor, every
// load and store of a warp is a contiguous run of 4-byte elements.  Block = 128 consecutive anchors of one level.
__global__ void __launch_bounds__(kDecThreads) k_decode_infer_any(LevelSet L, int nc, float *__restrict__ out) {
    const int b = blockIdx.y, chunk = blockIdx.x;
    const int tid = threadIdx.x;
    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && chunk >= L.chunk_off[k]) l = k;
    const int hw = L.hw[l], w = L.w[l];
    const float st = L.stride[l];
    const int c0 = (chunk - L.chunk_off[l]) * kChunk;
    const int i = c0 + tid;
    const bool live = i < hw;
    const int R = 4 + nc;
    const int A = L.A_true;
    const int a = L.a_true_off[l] + i;
    if (live) {
        const float *xin = L.x[l] + (size_t)b * L.C * hw + i;
        float dm[4];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            float z[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) z[j] = __ldg(xin + (size_t)(s * 16 + j) * hw);
            float m2, inv_s;
            dm[s] = dfl_stats16(z, m2, inv_s);
        }
        const int iy = i / w, ix = i - iy * w;
        const float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
        // explicit roundings: dm = d * inv_s must not be contracted into these (the other two decode kernels pass it
        // through shared memory; the three agree bit for bit)
        const float x1 = YS_SUB(gx, dm[0]), y1 = YS_SUB(gy, dm[1]), x2 = YS_ADD(gx, dm[2]), y2 = YS_ADD(gy, dm[3]);
        float o[4];
        o[0] = (x1 + x2) / 2 * st;      // dist2bbox(xywh=True) * strides  (models/yolo11_modules.py:263, :309-311)
        o[1] = (y1 + y2) / 2 * st;
        o[2] = (x2 - x1) * st;
        o[3] = (y2 - y1) * st;
#pragma unroll
        for (int c = 0; c < 4; ++c) out[((size_t)b * R + c) * A + a] = o[c];
        const float *xc = xin + (size_t)64 * hw;
        for (int c = 0; c < nc; ++c) out[((size_t)b * R + 4 + c) * A + a] = sigmoid_fast(__ldg(xc + (size_t)c * hw));
    }
}

// Anchor-major output (B, A, 4+nc) for ANY head shape -- the layout decode_predictions_for_metrics consumes
// (train_yolo11_cuda.py:289-301).  Block = 64 consecutive anchors of one level (half a 128-anchor chunk).
//   1. The block's whole input tile, [4*16 + nc channel rows][64 anchors], comes into shared memory by cp.async --
//      16-byte copies when the level's rows are 16-byte aligned, 4-byte copies otherwise -- so no register holds a load
//      and every byte of the tile is in flight at once; three blocks per SM keep ~110 KB in flight per SM.  (The
//      register-fed version before this one was bound by load latency: long scoreboard 12 per issue, 0.73 of the HBM
//      peak at 96 registers, 0.61 at 66.)
//   2. Decode out of shared memory, lane = anchor (conflict-free reads of a channel row), results staged as rows of
//      P floats:  VEC (P = R, R % 8 == 4, e.g. nc = 80): four channels of one anchor are one 16-byte word, and within a
//      quarter warp the words of consecutive anchors fall in distinct 16-byte bank groups (R / 4 is odd);
//      !VEC (P odd): scalar stores, conflict-free because P is odd.
//   3. The staged rows leave as one contiguous run (float4 when VEC).
constexpr int kAmCols = 64;
template <bool VEC>
__global__ void __launch_bounds__(kDecThreads) k_decode_infer_am(LevelSet L, int nc, int P, float *__restrict__ out) {
    extern __shared__ __align__(16) float s_dyn_dec[];
    __shared__ float s_d[4][kAmCols];
    const int b = blockIdx.y, chunk = blockIdx.x >> 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && chunk >= L.chunk_off[k]) l = k;
    const int hw = L.hw[l], w = L.w[l];
    const float st = L.stride[l];
    const int c0 = (chunk - L.chunk_off[l]) * kChunk + (blockIdx.x & 1) * kAmCols;
    const int rows = min(kAmCols, hw - c0);
    if (rows <= 0) return;                           // second half of a chunk that ends in its first half
    const int R = 4 + nc, C = 64 + nc;
    float *tile = s_dyn_dec;                         // [C][kAmCols]
    float *s_out = s_dyn_dec + (size_t)C * kAmCols;  // [kAmCols][P]
    const float *xin = L.x[l] + (size_t)b * L.C * hw + c0;

    // ---- 1. input tile -> shared memory
    if (L.tma[l]) {                                  // rows 16-byte aligned (h*w % 4 == 0 and aligned base): c0 % 64 == 0
        const int n16 = (rows + 3) >> 2;             // hw % 4 == 0, so a 16-byte piece never crosses the row's end
        for (int i = tid; i < C * 16; i += kDecThreads) {
            const int r = i >> 4, q = i & 15;
            if (q < n16) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(tile + r * kAmCols + q * 4);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(xin + (size_t)r * hw + q * 4) : "memory");
            }
        }
    } else {
        for (int i = tid; i < C * kAmCols; i += kDecThreads) {
            const int r = i >> 6, q = i & 63;
            if (q < rows) {
                const unsigned dst = (unsigned)__cvta_generic_to_shared(tile + r * kAmCols + q);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(xin + (size_t)r * hw + q) : "memory");
            }
        }
    }
    asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- 2a. box side `warp` of the 64 anchors (columns past `rows` hold stale shared memory: computed, never copied out)
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int r = lane + 32 * k;
        float z[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) z[j] = tile[(warp * 16 + j) * kAmCols + r];
        float m2, inv_s;
        s_d[warp][r] = dfl_stats16(z, m2, inv_s);
    }
    // ---- 2b. classes: a warp takes the groups of four channels g = warp, warp + 4, ...
    for (int c = warp * 4; c < nc; c += 16) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int r = lane + 32 * k;
            const float *tc = tile + (size_t)(64 + c) * kAmCols + r;
            float *dst = s_out + (size_t)r * P + 4 + c;
            if (VEC) {
                *reinterpret_cast<float4 *>(dst) = make_float4(sigmoid_fast(tc[0]), sigmoid_fast(tc[kAmCols]),
                                                                sigmoid_fast(tc[2 * kAmCols]), sigmoid_fast(tc[3 * kAmCols]));
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (c + j < nc) dst[j] = sigmoid_fast(tc[j * kAmCols]);
            }
        }
    }
    __syncthreads();
    if (tid < kAmCols) {   // box rows: thread = one anchor
        const int r = tid, i = c0 + r;
        const int iy = i / w, ix = i - iy * w;
        const float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
        const float x1 = gx - s_d[0][r], y1 = gy - s_d[1][r];
        const float x2 = gx + s_d[2][r], y2 = gy + s_d[3][r];
        // dist2bbox(xywh=True) * strides  (models/yolo11_modules.py:263, :309-311)
        const float4 o = make_float4((x1 + x2) / 2 * st, (y1 + y2) / 2 * st, (x2 - x1) * st, (y2 - y1) * st);
        float *dst = s_out + (size_t)r * P;
        if (VEC) *reinterpret_cast<float4 *>(dst) = o;
        else { dst[0] = o.x; dst[1] = o.y; dst[2] = o.z; dst[3] = o.w; }
    }
    __syncthreads();
    // ---- 3. copy-out
    float *dst = out + ((size_t)b * L.A_true + L.a_true_off[l] + c0) * R;
    const int total = rows * R;
    if (VEC) {
        const float4 *s4 = reinterpret_cast<const float4 *>(s_out);
        float4 *d4 = reinterpret_cast<float4 *>(dst);
#pragma unroll 4
        for (int i = tid; i < (total >> 2); i += kDecThreads) d4[i] = s4[i];
    } else {
        int row = tid / R, col = tid - row * R;            // kDecThreads = q R + rem: advance (row, col) without dividing
        const int q = kDecThreads / R, rem = kDecThreads - q * R;
        for (int i = tid; i < total; i += kDecThreads) {
            dst[i] = s_out[(size_t)row * P + col];
            row += q; col += rem;
            if (col >= R) { col -= R; ++row; }
        }
    }
}

// One thread = one row i and four consecutive columns j (a 16-byte store when the row pitch allows it).
__global__ void k_pairwise_iou(const float4 *__restrict__ b1, int N, const float4 *__restrict__ b2, int M, int kind,
                               float *__restrict__ out) {
    const int j0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int i = blockIdx.y;
    if (j0 >= M || i >= N) return;
    const float4 p = b1[i];
    const float area_p = kind == 0 ? area_plain(p) : area_eps_h(p.x, p.y, p.z, p.w);
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const float4 q = b2[min(j0 + u, M - 1)];
        if (kind == 0) v[u] = iou_nms(p, area_p, q, area_plain(q));
        else {
            // bbox_iou(xywh=False) without the assigner's clamp (losses/yolo_v8_loss.py:31-44)
            const float inter = inter_area(p.x, p.y, p.z, p.w, q.x, q.y, q.z, q.w);
            const float uni = YS_ADD(YS_SUB(YS_ADD(area_p, area_eps_h(q.x, q.y, q.z, q.w)), inter), kEpsIou);
            v[u] = YS_DIV(inter, uni);
        }
    }
    float *row = out + (size_t)i * M + j0;
    if (j0 + 3 < M && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
        *reinterpret_cast<float4 *>(row) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (j0 + u < M) row[u] = v[u];
    }
}

// Elementwise bbox_iou, all modes (losses/yolo_v8_loss.py:12-61), op order as the reference.
__global__ void k_bbox_iou(const float4 *__restrict__ box1, const float4 *__restrict__ box2, int n, int xywh, int mode,
                           float eps, float *__restrict__ out, const float *__restrict__ gout, float4 *__restrict__ grad1) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = box1[i], b = box2[i];
    float ax1, ay1, ax2, ay2, bx1, by1, bx2, by2, w1, h1, w2, h2;
    if (xywh) {
        w1 = a.z; h1 = a.w; w2 = b.z; h2 = b.w;
        const float hw1 = YS_DIV(w1, 2.f), hh1 = YS_DIV(h1, 2.f), hw2 = YS_DIV(w2, 2.f), hh2 = YS_DIV(h2, 2.f);
        ax1 = YS_SUB(a.x, hw1); ax2 = YS_ADD(a.x, hw1); ay1 = YS_SUB(a.y, hh1); ay2 = YS_ADD(a.y, hh1);
        bx1 = YS_SUB(b.x, hw2); bx2 = YS_ADD(b.x, hw2); by1 = YS_SUB(b.y, hh2); by2 = YS_ADD(b.y, hh2);
    } else {
        ax1 = a.x; ay1 = a.y; ax2 = a.z; ay2 = a.w; bx1 = b.x; by1 = b.y; bx2 = b.z; by2 = b.w;
        w1 = YS_SUB(ax2, ax1); h1 = YS_ADD(YS_SUB(ay2, ay1), eps);
        w2 = YS_SUB(bx2, bx1); h2 = YS_ADD(YS_SUB(by2, by1), eps);
    }
    const float inter = inter_area(ax1, ay1, ax2, ay2, bx1, by1, bx2, by2);
    const float uni = YS_ADD(YS_SUB(YS_ADD(YS_MUL(w1, h1), YS_MUL(w2, h2)), inter), eps);
    const float iou = YS_DIV(inter, uni);
    float r = iou;
    if (mode != 0) {
        const float cw = YS_SUB(fmaxf2(ax2, bx2), fminf2(ax1, bx1));
        const float ch = YS_SUB(fmaxf2(ay2, by2), fminf2(ay1, by1));
        if (mode == 2 || mode == 3) {
            const float c2 = YS_ADD(YS_ADD(YS_MUL(cw, cw), YS_MUL(ch, ch)), eps);
            const float sx = YS_SUB(YS_SUB(YS_ADD(bx1, bx2), ax1), ax2), sy = YS_SUB(YS_SUB(YS_ADD(by1, by2), ay1), ay2);
            const float rho2 = YS_DIV(YS_ADD(YS_MUL(sx, sx), YS_MUL(sy, sy)), 4.f);
            if (mode == 3) {
                const float dat = YS_SUB(atanf(YS_DIV(w2, h2)), atanf(YS_DIV(w1, h1)));
                const float v = YS_MUL(0.40528473456935108578f, YS_MUL(dat, dat));
                const float alpha = YS_DIV(v, YS_ADD(YS_SUB(v, iou), YS_ADD(1.f, eps)));
                r = YS_SUB(iou, YS_ADD(YS_DIV(rho2, c2), YS_MUL(v, alpha)));
            } else {
                r = YS_SUB(iou, YS_DIV(rho2, c2));
            }
        } else {
            const float c_area = YS_ADD(YS_MUL(cw, ch), eps);
            r = YS_SUB(iou, YS_DIV(YS_SUB(c_area, uni), c_area));
        }
    }
    out[i] = r;
    if (grad1) {
        // supported: xyxy, mode 3 (CIoU), eps = 1e-7  -- the BboxLoss use (losses/yolo_v8_loss.py:299)
        float g[4];
        (void)ciou_with_grad(make_float4(ax1, ay1, ax2, ay2), make_float4(bx1, by1, bx2, by2), g);
        const float go = gout ? gout[i] : 1.f;
        grad1[i] = make_float4(g[0] * go, g[1] * go, g[2] * go, g[3] * go);
    }
}

}  // namespace ys

using namespace ys;

// build_levels lives in ys_loss.cu; re-declare a light version here for the decode entry point
static int build_levels_dec(LevelSet &L, int n_levels, const float *const *xs, const int *hs, const int *ws,
                            const float *strides, int nc, int reg_max, bool &all_vec) {
    YS_REQUIRE(n_levels >= 1 && n_levels <= YS_MAX_LEVELS, YS_ERR_BAD_ARG, "n_levels must be in [1,%d]", YS_MAX_LEVELS);
    YS_REQUIRE(reg_max == 16, YS_ERR_UNSUPPORTED, "reg_max must be 16");
    YS_REQUIRE(nc >= 1, YS_ERR_BAD_ARG, "nc must be >= 1");
    memset(&L, 0, sizeof(L));
    L.n = n_levels;
    L.C = 4 * reg_max + nc;
    int a = 0, ch = 0;
    all_vec = true;
    for (int i = 0; i < n_levels; ++i) {
        YS_REQUIRE(xs[i] != nullptr && hs[i] > 0 && ws[i] > 0, YS_ERR_BAD_ARG, "level %d: null pointer or empty shape", i);
        YS_REQUIRE(((uintptr_t)xs[i] & 3) == 0, YS_ERR_BAD_ARG, "level %d: pointer must be 4-byte aligned", i);
        L.x[i] = xs[i]; L.g[i] = nullptr; L.w[i] = ws[i]; L.hw[i] = hs[i] * ws[i]; L.stride[i] = strides[i];
        L.tma[i] = (L.hw[i] % 4 == 0) && (((uintptr_t)xs[i] & 15) == 0);
        all_vec = all_vec && L.tma[i];
        L.a_off[i] = a; L.a_true_off[i] = a; L.chunk_off[i] = ch;      // the decode writes the reference's dense numbering only
        a += L.hw[i];
        ch += (L.hw[i] + kChunk - 1) / kChunk;
    }
    for (int i = n_levels; i <= YS_MAX_LEVELS; ++i) { L.a_off[i] = a; L.a_true_off[i] = a; L.chunk_off[i] = ch; }
    for (int i = n_levels; i < YS_MAX_LEVELS; ++i) { L.w[i] = 1; L.stride[i] = 1.f; }
    L.A = a;
    L.A_true = a;
    L.n_chunks = ch;
    return YS_OK;
}

extern "C" int ys_decode_infer(int n_levels, const float *const *level_x_host, const int *level_h_host,
                               const int *level_w_host, const float *level_stride_host, int B, int nc, int reg_max,
                               float *out, int layout, void *stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    YS_REQUIRE(out != nullptr && ((uintptr_t)out & 15) == 0, YS_ERR_BAD_ARG, "out must be non-null and 16-byte aligned");
    YS_REQUIRE(B >= 1 && B <= 65535, YS_ERR_BAD_ARG, "B must be in [1,65535]");
    YS_REQUIRE(layout == 0 || layout == 1, YS_ERR_BAD_ARG, "layout must be 0 or 1");
    LevelSet L;
    bool all_vec = true;
    int rc = build_levels_dec(L, n_levels, level_x_host, level_h_host, level_w_host, level_stride_host, nc, reg_max, all_vec);
    if (rc) return rc;
    dim3 grid(L.n_chunks, B);
    if (layout == 0) {
        // float4 kernel when every level's rows and every output row are 16-byte aligned
        if (all_vec && L.A % 4 == 0) k_decode_infer<<<grid, kDecThreads, 0, stream>>>(L, nc, out);
        else k_decode_infer_any<<<grid, kDecThreads, 0, stream>>>(L, nc, out);
    } else {
        const int R = 4 + nc;
        const bool vec = (R & 7) == 4;                 // rows of R floats: 16-byte words, R / 4 odd
        const int P = vec ? R : (R | 1);
        const size_t smem = ((size_t)(64 + nc) * kAmCols + (size_t)kAmCols * P) * sizeof(float);
        YS_REQUIRE(smem <= 200 * 1024, YS_ERR_UNSUPPORTED, "nc too large for the staged layout (%d)", nc);
        static DeviceOnce once_dec[2];
        rc = optin_smem(vec ? (const void *)k_decode_infer_am<true> : (const void *)k_decode_infer_am<false>, 200 * 1024,
                        once_dec[vec ? 0 : 1], "k_decode_infer_am");
        if (rc) return rc;
        dim3 grid_am(2 * L.n_chunks, B);
        if (vec) k_decode_infer_am<true><<<grid_am, kDecThreads, smem, stream>>>(L, nc, P, out);
        else k_decode_infer_am<false><<<grid_am, kDecThreads, smem, stream>>>(L, nc, P, out);
    }
    YS_CHECK_LAUNCH("k_decode_infer");
    return YS_OK;
}

extern "C" int ys_pairwise_iou(const float *boxes1, int N, const float *boxes2, int M, int kind, float *out, void *stream_v) {
    YS_REQUIRE(N >= 0 && M >= 0, YS_ERR_BAD_ARG, "negative size");
    if (N == 0 || M == 0) return YS_OK;
    YS_REQUIRE(boxes1 && boxes2 && out, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE((((uintptr_t)boxes1 | (uintptr_t)boxes2) & 15) == 0, YS_ERR_BAD_ARG, "box pointers must be 16-byte aligned");
    YS_REQUIRE(kind == 0 || kind == 1, YS_ERR_BAD_ARG, "kind must be 0 or 1");
    YS_REQUIRE(N <= 65535, YS_ERR_UNSUPPORTED, "N must be <= 65535");
    dim3 grid((M + 1023) / 1024, N);
    k_pairwise_iou<<<grid, 256, 0, (cudaStream_t)stream_v>>>((const float4 *)boxes1, N, (const float4 *)boxes2, M, kind, out);
    YS_CHECK_LAUNCH("k_pairwise_iou");
    return YS_OK;
}

extern "C" int ys_bbox_iou(const float *box1, const float *box2, int n, int xywh, int mode, float eps, float *out,
                           const float *gout, float *grad_box1, void *stream_v) {
    YS_REQUIRE(n >= 0, YS_ERR_BAD_ARG, "negative size");
    if (n == 0) return YS_OK;
    YS_REQUIRE(box1 && box2 && out, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE((((uintptr_t)box1 | (uintptr_t)box2 | (uintptr_t)grad_box1) & 15) == 0, YS_ERR_BAD_ARG, "box pointers must be 16-byte aligned");
    YS_REQUIRE(mode >= 0 && mode <= 3, YS_ERR_BAD_ARG, "mode must be 0..3");
    if (grad_box1) YS_REQUIRE(!xywh && mode == 3 && eps == 1e-7f, YS_ERR_UNSUPPORTED, "gradient is implemented for xyxy CIoU with eps=1e-7 only");
    k_bbox_iou<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream_v>>>((const float4 *)box1, (const float4 *)box2, n, xywh, mode, eps, out,
                                                                   gout, (float4 *)grad_box1);
    YS_CHECK_LAUNCH("k_bbox_iou");
    return YS_OK;
}
