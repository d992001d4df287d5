// ys_loss.cu -- fused detection loss (assign + BCE + CIoU + DFL, forward and backward) and the dense
// TaskAlignedAssigner entry point.  Replaces losses/yolo_v8_loss.py of the reference; see include/ys_b200.h.
//
// Launch sequence of ys_detection_loss (one stream, no host sync):
//   k_pack_gt        flat GT list -> per-image CSR (stable), pixel boxes        [replaces preprocess :501-527]
//   k_assign_scan    decode + pairwise IoU / in-box + preliminary assignment    [reads the 64 box channels]
//   k_assign_fixup   forced assignments of the two reference loops, tss         [tiny]
//   k_loss_fwd_bwd   BCE + CIoU + DFL forward and backward, NCHW in, NCHW out   [reads 144 ch, writes 144 ch]
//   k_loss_finalize  ordered fp64 reduction of the per-block partial sums        [tiny]
#include "ys_assign.cuh"

namespace ys {

// =====================================================================================================
// GT packing
// =====================================================================================================
constexpr int kPackThreads = 1024;
constexpr int kPackSmemB = 8192;

__global__ void __launch_bounds__(kPackThreads)
k_pack_gt(const long long *__restrict__ batch_idx, const long long *__restrict__ cls, const float4 *__restrict__ bboxes,
          int N, int B, int nc, float sx, float sy, int *__restrict__ gt_off, int *__restrict__ rank,
          int *__restrict__ cursor_g, float4 *__restrict__ gt_box, int *__restrict__ gt_label,
          unsigned long long *__restrict__ best_key, int *__restrict__ owned_cnt, int *__restrict__ gt_flags) {
    __shared__ int s_cursor[kPackSmemB];
    __shared__ int s_part[kPackThreads / 32];
    __shared__ int s_unsorted;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    volatile int *cursor = (B <= kPackSmemB) ? (volatile int *)s_cursor : (volatile int *)cursor_g;
    for (int i = tid; i < B; i += kPackThreads) cursor[i] = 0;
    if (tid == 0) s_unsorted = 0;
    __syncthreads();
    // sorted and all ids valid?  (what the reference collate produces)
    for (int i = tid; i < N; i += kPackThreads) {
        long long id = batch_idx[i];
        bool bad = id < 0 || id >= B || (i > 0 && batch_idx[i - 1] > id);
        if (bad) s_unsorted = 1;
    }
    __syncthreads();
    if (!s_unsorted) {
        // counts by segment boundaries
        for (int i = tid; i < N; i += kPackThreads) {
            int id = (int)batch_idx[i];
            if (i == 0 || (int)batch_idx[i - 1] != id) {
                // first row of image id; its count = next boundary - i, found by scanning forward is O(N);
                // instead record the start and derive counts below
                cursor[id] = -(i + 1);   // mark start (negative, 1-based)
            }
            rank[i] = i;   // temporarily the absolute position; made relative below
        }
        __syncthreads();
        // start[id] for present ids; counts come from a backward fill by one warp-strided pass
        if (tid == 0) {
            int next = N;
            for (int id = B - 1; id >= 0; --id) {
                int c = cursor[id];
                if (c < 0) {
                    int start = -c - 1;
                    cursor[id] = next - start;
                    next = start;
                } else {
                    cursor[id] = 0;
                }
            }
        }
        __syncthreads();
    } else if (warp == 0) {
        // general case: stable rank within image, one warp walking the list 32 rows at a time
        for (int base = 0; base < N; base += 32) {
            int i = base + lane;
            long long id = i < N ? batch_idx[i] : -1;
            bool ok = id >= 0 && id < B;
            int key = ok ? (int)id : -1 - lane;
            unsigned peers = __match_any_sync(0xffffffffu, key);
            int r = __popc(peers & ((1u << lane) - 1u));
            int basecnt = ok ? cursor[(int)id] : 0;
            __syncwarp();
            if (ok) {
                rank[i] = basecnt + r;
                if (r == 0) cursor[(int)id] = basecnt + __popc(peers);
            } else if (i < N) {
                rank[i] = -1;
            }
            __syncwarp();
        }
    }
    __syncthreads();
    // exclusive scan of the counts -> gt_off (each thread owns a contiguous slice)
    const int per = (B + kPackThreads - 1) / kPackThreads;
    const int lo = min(B, tid * per), hi = min(B, lo + per);
    int local = 0;
    for (int i = lo; i < hi; ++i) local += cursor[i];
    int incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_part[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int v = s_part[lane];
        int inc2 = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, inc2, o);
            if (lane >= o) inc2 += t;
        }
        s_part[lane] = inc2 - v;
    }
    __syncthreads();
    int run = s_part[warp] + incl - local;
    for (int i = lo; i < hi; ++i) {
        gt_off[i] = run;
        run += cursor[i];
    }
    if (hi == B && lo < B) gt_off[B] = run;
    if (B == 0 && tid == 0) gt_off[0] = 0;
    __syncthreads();
    __threadfence_block();
    // placement
    const bool sorted = !s_unsorted;
    for (int i = tid; i < N; i += kPackThreads) {
        best_key[i] = 0ull;
        owned_cnt[i] = 0;
        gt_flags[i] = 0;
        long long id = batch_idx[i];
        if (id < 0 || id >= B) continue;
        int dst = sorted ? i : (gt_off[(int)id] + rank[i]);
        float4 q = bboxes[i];
        // targets_bbox * [sx, sy, sx, sy]   (losses/yolo_v8_loss.py:515-516)
        gt_box[dst] = make_float4(YS_MUL(q.x, sx), YS_MUL(q.y, sy), YS_MUL(q.z, sx), YS_MUL(q.w, sy));
        long long c = cls[i];
        c = c < 0 ? 0 : (c > nc - 1 ? nc - 1 : c);
        gt_label[dst] = (int)c;
    }
}

// =====================================================================================================
// Loss forward + backward
// =====================================================================================================
constexpr int kLossThreads = 128;   // 4 warps: warp w owns box side w (16 channels) and a quarter of the classes

struct LossParams {
    const int *tgi;          // (B,A) -1 = background
    const float *norm;       // (B,A)
    const int *gt_off;       // (B+1)
    const float4 *gt_box;    // px
    const int *gt_label;
    const float *tss;        // [0]
    float *partials;         // (gridDim.x*gridDim.y, 4): box, cls, dfl partial sums (unnormalised)
    int nc;
    float hyp_box, hyp_cls, hyp_dfl;
    float batch;             // (float)B
};

template <bool WRITE_GRAD>
__global__ void __launch_bounds__(kLossThreads) k_loss_fwd_bwd(LevelSet L, LossParams P) {
    __shared__ float s_d[4][kChunk];       // ltrb expectation per side per anchor of the chunk
    __shared__ float s_sum[4][3];

    const int b = blockIdx.y;
    const int chunk = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && chunk >= L.chunk_off[k]) l = k;
    const int hw = L.hw[l], w = L.w[l];
    const float st = L.stride[l];
    const int i0 = (chunk - L.chunk_off[l]) * kChunk + lane * 4;   // first of this thread's 4 anchors (within level)
    const bool live = i0 < hw;                                     // hw % 4 == 0 => all 4 anchors live together
    const size_t img = (size_t)b * L.C * hw;
    const float *xin = L.x[l] + img + i0;
    float *gout = WRITE_GRAD ? (L.g[l] + img + i0) : nullptr;
    const int a0 = L.a_off[l] + i0;

    const float tss = P.tss[0];
    const float inv_tss = 1.0f / tss;

    int tg[4] = {-1, -1, -1, -1};
    float nr[4] = {0.f, 0.f, 0.f, 0.f};
    int lab[4] = {-1, -1, -1, -1};
    if (live) {
        int4 t4 = *reinterpret_cast<const int4 *>(P.tgi + (size_t)b * L.A + a0);
        float4 n4 = *reinterpret_cast<const float4 *>(P.norm + (size_t)b * L.A + a0);
        tg[0] = t4.x; tg[1] = t4.y; tg[2] = t4.z; tg[3] = t4.w;
        nr[0] = n4.x; nr[1] = n4.y; nr[2] = n4.z; nr[3] = n4.w;
    }
    const int g0 = P.gt_off[b];
    const bool any_fg = (tg[0] >= 0) | (tg[1] >= 0) | (tg[2] >= 0) | (tg[3] >= 0);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (tg[k] >= 0) lab[k] = P.gt_label[g0 + tg[k]];

    float acc_box = 0.f, acc_cls = 0.f, acc_dfl = 0.f;

    // ------------------------------------------------------------------ box side `warp`
    // z[j][k]: logit of bin j for anchor k; becomes the softmax probability in place.
    float4 z[16];
    float dmine[4] = {0.f, 0.f, 0.f, 0.f}, lse[4] = {0.f, 0.f, 0.f, 0.f};
    const int cbox = warp * 16;
    if (any_fg) {
#pragma unroll
        for (int j = 0; j < 16; ++j) z[j] = __ldg(reinterpret_cast<const float4 *>(xin + (size_t)(cbox + j) * hw));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float zz[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) zz[j] = (&z[j].x)[k];
            float m, es;
            dmine[k] = dfl_expect16(zz, m, es);
            lse[k] = m + logf(es);
            // keep raw logits in z for the cross-entropy lookups; probabilities are recomputed below
        }
    }
    *reinterpret_cast<float4 *>(&s_d[warp][lane * 4]) = make_float4(dmine[0], dmine[1], dmine[2], dmine[3]);
    __syncthreads();

    if (any_fg) {
        float4 d0 = *reinterpret_cast<const float4 *>(&s_d[0][lane * 4]);
        float4 d1 = *reinterpret_cast<const float4 *>(&s_d[1][lane * 4]);
        float4 d2 = *reinterpret_cast<const float4 *>(&s_d[2][lane * 4]);
        float4 d3 = *reinterpret_cast<const float4 *>(&s_d[3][lane * 4]);
        float gd[4];      // d(total)/d(dmine[k]) from the CIoU term
        float cdfl[4];    // DFL gradient scale per anchor
        int tl[4];
        float wl[4], wr[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            gd[k] = 0.f; cdfl[k] = 0.f; tl[k] = 0; wl[k] = 0.f; wr[k] = 0.f;
            if (tg[k] < 0) continue;
            const int i = i0 + k;
            const int iy = i / w, ix = i - iy * w;
            const float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
            const float dd0 = (&d0.x)[k], dd1 = (&d1.x)[k], dd2 = (&d2.x)[k], dd3 = (&d3.x)[k];
            const float4 pb = make_float4(gx - dd0, gy - dd1, gx + dd2, gy + dd3);       // grid units
            const float4 q = __ldg(P.gt_box + g0 + tg[k]);
            const float4 tb = make_float4(q.x / st, q.y / st, q.z / st, q.w / st);      // target_bboxes /= stride (:479)
            float g4[4];
            const float ciou = ciou_with_grad(pb, tb, g4);
            const float wgt = nr[k];                                                    // target_scores.sum(-1) (:298)
            if (warp == 0) acc_box += (1.0f - ciou) * wgt;
            // d/d(dist): x1 = gx - d0 -> -g4[0]; y1 -> -g4[1]; x2 = gx + d2 -> g4[2]; y2 -> g4[3]
            const float dsign = (warp == 0) ? -g4[0] : (warp == 1) ? -g4[1] : (warp == 2) ? g4[2] : g4[3];
            gd[k] = -(P.hyp_box * P.batch * inv_tss) * wgt * dsign;                     // L = (1-ciou)*w/tss
            // DFL target for my side (:304, :327-330, :315)
            float t = (warp == 0) ? gx - tb.x : (warp == 1) ? gy - tb.y : (warp == 2) ? tb.z - gx : tb.w - gy;
            t = fminf(fmaxf(t, 0.f), 14.99f);   // reg_max - 1 - 0.01
            const int tli = (int)t;
            tl[k] = tli;
            wl[k] = (float)(tli + 1) - t;
            wr[k] = 1.0f - wl[k];
            float zl = 0.f, zr = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float v = (&z[j].x)[k];
                zl = (j == tli) ? v : zl;
                zr = (j == tli + 1) ? v : zr;
            }
            const float ce = (lse[k] - zl) * wl[k] + (lse[k] - zr) * wr[k];
            acc_dfl += ce * 0.25f * wgt;                                                // mean over 4 sides (:324) * weight
            cdfl[k] = P.hyp_dfl * P.batch * inv_tss * 0.25f * wgt;
        }
        if (WRITE_GRAD) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                float4 o;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float gval = 0.f;
                    if (tg[k] >= 0) {
                        const float p = expf((&z[j].x)[k] - lse[k]);
                        gval = gd[k] * p * ((float)j - dmine[k]) +
                               cdfl[k] * (p * (wl[k] + wr[k]) - (j == tl[k] ? wl[k] : 0.f) - (j == tl[k] + 1 ? wr[k] : 0.f));
                    }
                    (&o.x)[k] = gval;
                }
                *reinterpret_cast<float4 *>(gout + (size_t)(cbox + j) * hw) = o;
            }
        }
    } else if (WRITE_GRAD && live) {
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = 0; j < 16; ++j) *reinterpret_cast<float4 *>(gout + (size_t)(cbox + j) * hw) = zero;
    }

    // ------------------------------------------------------------------ class channels of this warp
    if (live) {
        const int per = (P.nc + 3) >> 2;
        const int c_lo = warp * per, c_hi = min(P.nc, c_lo + per);
        const float gscale = P.hyp_cls * P.batch * inv_tss;
        const float *xc = xin + (size_t)64 * hw;
        float *gc = WRITE_GRAD ? gout + (size_t)64 * hw : nullptr;
#pragma unroll 4
        for (int c = c_lo; c < c_hi; ++c) {
            const float4 xv = __ldg(reinterpret_cast<const float4 *>(xc + (size_t)c * hw));
            float4 o;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float t = (c == lab[k]) ? nr[k] : 0.f;
                float ls, sg;
                bce_logits((&xv.x)[k], t, ls, sg);
                acc_cls += ls;
                (&o.x)[k] = (sg - t) * gscale;
            }
            if (WRITE_GRAD) *reinterpret_cast<float4 *>(gc + (size_t)c * hw) = o;
        }
    }

    // ------------------------------------------------------------------ block partial sums (fixed order)
    acc_box = warp_sum(acc_box);
    acc_cls = warp_sum(acc_cls);
    acc_dfl = warp_sum(acc_dfl);
    if (lane == 0) { s_sum[warp][0] = acc_box; s_sum[warp][1] = acc_cls; s_sum[warp][2] = acc_dfl; }
    __syncthreads();
    if (tid == 0) {
        float *o = P.partials + ((size_t)b * gridDim.x + chunk) * 4;
        o[0] = s_sum[0][0] + s_sum[1][0] + s_sum[2][0] + s_sum[3][0];
        o[1] = s_sum[0][1] + s_sum[1][1] + s_sum[2][1] + s_sum[3][1];
        o[2] = s_sum[0][2] + s_sum[1][2] + s_sum[2][2] + s_sum[3][2];
        o[3] = 0.f;
    }
}

__global__ void __launch_bounds__(1024)
k_loss_finalize(const float *__restrict__ partials, int n, const float *__restrict__ tss, float hyp_box, float hyp_cls,
                float hyp_dfl, float batch, float *__restrict__ loss_out) {
    __shared__ double s[32][3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double a0 = 0, a1 = 0, a2 = 0;
    for (int i = tid; i < n; i += 1024) {
        const float4 v = reinterpret_cast<const float4 *>(partials)[i];
        a0 += v.x; a1 += v.y; a2 += v.z;
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
    if (lane == 0) { s[warp][0] = a0; s[warp][1] = a1; s[warp][2] = a2; }
    __syncthreads();
    if (tid == 0) {
        double t0 = 0, t1 = 0, t2 = 0;
        for (int w = 0; w < 32; ++w) { t0 += s[w][0]; t1 += s[w][1]; t2 += s[w][2]; }
        const double inv = 1.0 / (double)tss[0];
        // loss[0] = box/tss*hyp_box, loss[1] = cls/tss*hyp_cls, loss[2] = dfl/tss*hyp_dfl (:475, :300, :306, :495-497)
        const float lb = (float)(t0 * inv) * hyp_box;
        const float lc = (float)(t1 * inv) * hyp_cls;
        const float ld = (float)(t2 * inv) * hyp_dfl;
        loss_out[0] = (lb + lc + ld) * batch;   // loss.sum() * batch_size (:499)
        loss_out[1] = lb;
        loss_out[2] = lc;
        loss_out[3] = ld;
    }
}

__global__ void k_scale_inplace(float *__restrict__ data, size_t n, const float *__restrict__ scale) {
    const float s = *scale;
    if (s == 1.0f) return;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) data[i] *= s;
}

// Dense assigner outputs (losses/yolo_v8_loss.py:246-270): labels / boxes for ALL anchors (background gets
// GT slot 0), one-hot * norm scores, bool fg, i64 gt index.
__global__ void k_assign_emit(int B, int A, int G, int nc, const int *__restrict__ tgi, const float *__restrict__ norm,
                              const float *__restrict__ gt_labels, const float4 *__restrict__ gt_bboxes,
                              float *__restrict__ target_labels, float4 *__restrict__ target_bboxes,
                              float *__restrict__ target_scores, unsigned char *__restrict__ fg_mask,
                              long long *__restrict__ target_gt_idx) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)B * A) return;
    const int b = (int)(i / A);
    const int g = tgi[i];
    const int slot = g < 0 ? 0 : g;
    float lab = gt_labels[(size_t)b * G + slot];
    lab = lab < 0.f ? 0.f : (lab > (float)nc ? (float)nc : lab);       // clamp_(0, num_classes) (:256)
    target_labels[i] = lab;
    target_bboxes[i] = gt_bboxes[(size_t)b * G + slot];
    fg_mask[i] = g >= 0;
    target_gt_idx[i] = slot;
    if (target_scores) {
        float *row = target_scores + i * nc;
        const int li = (int)lab;
        const float v = g >= 0 ? norm[i] : 0.f;
        for (int c = 0; c < nc; ++c) row[c] = (c == li) ? v : 0.f;
    }
}

static int build_levels(LevelSet &L, int n_levels, const float *const *xs, float *const *gs, const int *hs,
                        const int *ws, const float *strides, int nc, int reg_max) {
    YS_REQUIRE(n_levels >= 1 && n_levels <= YS_MAX_LEVELS, YS_ERR_BAD_ARG, "n_levels must be in [1,%d]", YS_MAX_LEVELS);
    YS_REQUIRE(reg_max == 16, YS_ERR_UNSUPPORTED, "reg_max must be 16 (reference Detect fixes it, models/yolo11_modules.py:214)");
    YS_REQUIRE(nc >= 1, YS_ERR_BAD_ARG, "nc must be >= 1");
    L.n = n_levels;
    L.C = 4 * reg_max + nc;
    int a = 0, ch = 0;
    for (int i = 0; i < n_levels; ++i) {
        YS_REQUIRE(xs[i] != nullptr && hs[i] > 0 && ws[i] > 0, YS_ERR_BAD_ARG, "level %d: null pointer or empty shape", i);
        YS_REQUIRE(((uintptr_t)xs[i] & 15) == 0, YS_ERR_BAD_ARG, "level %d: pointer must be 16-byte aligned", i);
        YS_REQUIRE((hs[i] * ws[i]) % 4 == 0, YS_ERR_UNSUPPORTED, "level %d: h*w must be a multiple of 4 (float4 path)", i);
        L.x[i] = xs[i];
        L.g[i] = gs ? gs[i] : nullptr;
        if (gs) YS_REQUIRE(gs[i] != nullptr && ((uintptr_t)gs[i] & 15) == 0, YS_ERR_BAD_ARG, "level %d: bad grad pointer", i);
        L.w[i] = ws[i];
        L.hw[i] = hs[i] * ws[i];
        L.stride[i] = strides[i];
        L.a_off[i] = a;
        L.chunk_off[i] = ch;
        a += L.hw[i];
        ch += (L.hw[i] + kChunk - 1) / kChunk;
    }
    for (int i = n_levels; i <= YS_MAX_LEVELS; ++i) { L.a_off[i] = a; L.chunk_off[i] = ch; }
    for (int i = n_levels; i < YS_MAX_LEVELS; ++i) { L.x[i] = nullptr; L.g[i] = nullptr; L.w[i] = 1; L.hw[i] = 0; L.stride[i] = 1.f; }
    L.A = a;
    L.n_chunks = ch;
    return YS_OK;
}

struct LossWorkspace {
    int *gt_off, *rank, *cursor, *gt_label, *owned_cnt, *gt_flags, *forced, *tgi;
    float4 *gt_box, *pbox;
    unsigned long long *best_key;
    float *norm, *tss, *partials;
    double *img_sum;
    unsigned *done_counter;
    size_t bytes;
};

static void carve_loss(LossWorkspace &W, void *base, size_t cap, int B, int A, int n_gt) {
    Carver c(base, cap);
    const size_t N = (size_t)(n_gt > 0 ? n_gt : 1);
    const size_t chunks = (size_t)B * ((size_t)(A + kChunk - 1) / kChunk + YS_MAX_LEVELS);
    W.done_counter = c.take<unsigned>(64);        // zeroed once by the caller's first memset; self-resetting
    W.tss = c.take<float>(64);
    W.gt_off = c.take<int>((size_t)B + 1);
    W.cursor = c.take<int>((size_t)B + 1);
    W.rank = c.take<int>(N);
    W.gt_label = c.take<int>(N);
    W.owned_cnt = c.take<int>(N);
    W.gt_flags = c.take<int>(N);
    W.forced = c.take<int>(2 * N);
    W.gt_box = c.take<float4>(N);
    W.best_key = c.take<unsigned long long>(N);
    W.img_sum = c.take<double>((size_t)B);
    W.tgi = c.take<int>((size_t)B * A);
    W.norm = c.take<float>((size_t)B * A);
    W.pbox = c.take<float4>((size_t)B * A);
    W.partials = c.take<float>(chunks * 4);
    W.bytes = align_up(c.off, 256);
}

}  // namespace ys

using namespace ys;

extern "C" size_t ys_detection_loss_workspace_bytes(int B, int A, int n_gt) {
    LossWorkspace W;
    carve_loss(W, nullptr, 0, B > 0 ? B : 1, A > 0 ? A : 1, n_gt);
    return W.bytes;
}

extern "C" int ys_detection_loss(int n_levels, const float *const *level_x_host, float *const *level_grad_host,
                                 const int *level_h_host, const int *level_w_host, const float *level_stride_host,
                                 int B, int nc, int reg_max, const int64_t *batch_idx, const int64_t *cls,
                                 const float *bboxes, int n_gt, float scale_x, float scale_y, float hyp_box,
                                 float hyp_cls, float hyp_dfl, float alpha, float beta, float *loss_out,
                                 float *pred_boxes_px, int32_t *target_gt_idx, float *norm_out, int stages,
                                 void *workspace, size_t workspace_bytes, void *stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    YS_REQUIRE(B >= 1, YS_ERR_BAD_ARG, "B must be >= 1");
    YS_REQUIRE(n_gt >= 0, YS_ERR_BAD_ARG, "n_gt must be >= 0");
    YS_REQUIRE(loss_out != nullptr && workspace != nullptr, YS_ERR_BAD_ARG, "loss_out / workspace is null");
    YS_REQUIRE(n_gt == 0 || (batch_idx && cls && bboxes), YS_ERR_BAD_ARG, "GT pointers are null");
    YS_REQUIRE(n_gt == 0 || ((uintptr_t)bboxes & 15) == 0, YS_ERR_BAD_ARG, "bboxes must be 16-byte aligned");
    YS_REQUIRE(B <= 65535, YS_ERR_UNSUPPORTED, "B must be <= 65535 (grid.y)");
    LevelSet L;
    int rc = build_levels(L, n_levels, level_x_host, level_grad_host, level_h_host, level_w_host, level_stride_host, nc, reg_max);
    if (rc) return rc;
    LossWorkspace W;
    carve_loss(W, workspace, workspace_bytes, B, L.A, n_gt);
    YS_REQUIRE(W.bytes <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu", W.bytes, workspace_bytes);
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    if (pred_boxes_px) { YS_REQUIRE(((uintptr_t)pred_boxes_px & 15) == 0, YS_ERR_BAD_ARG, "pred_boxes_px must be 16-byte aligned"); }
    if (target_gt_idx) { YS_REQUIRE(((uintptr_t)target_gt_idx & 15) == 0, YS_ERR_BAD_ARG, "target_gt_idx must be 16-byte aligned"); }
    if (norm_out) { YS_REQUIRE(((uintptr_t)norm_out & 15) == 0, YS_ERR_BAD_ARG, "norm must be 16-byte aligned"); }

    if (stages == 0) stages = YS_STAGE_ALL;
    if (stages & YS_STAGE_PACK) {
        cudaMemsetAsync(W.done_counter, 0, sizeof(unsigned), stream);
        k_pack_gt<<<1, kPackThreads, 0, stream>>>((const long long *)batch_idx, (const long long *)cls,
                                                  (const float4 *)bboxes, n_gt, B, nc, scale_x, scale_y, W.gt_off, W.rank,
                                                  W.cursor, W.gt_box, W.gt_label, W.best_key, W.owned_cnt, W.gt_flags);
        YS_CHECK_LAUNCH("k_pack_gt");
    }

    AssignIO io;
    io.gt_off = W.gt_off; io.G = 0; io.gt_box = W.gt_box; io.gt_label_i = W.gt_label; io.gt_label_f = nullptr;
    io.gt_valid = nullptr; io.pd_box = nullptr; io.pd_scores = nullptr; io.anc = nullptr;
    io.best_key = W.best_key; io.owned_cnt = W.owned_cnt; io.gt_flags = W.gt_flags; io.forced = W.forced;
    io.pbox = pred_boxes_px ? (float4 *)pred_boxes_px : W.pbox;
    io.tgi = target_gt_idx ? (int *)target_gt_idx : W.tgi;
    io.norm = norm_out ? norm_out : W.norm;
    io.img_sum = W.img_sum; io.tss = W.tss; io.done_counter = W.done_counter;
    io.nc = nc; io.alpha = alpha; io.beta = beta; io.eps = 1e-9f;

    dim3 gs((L.A + kScanThreads - 1) / kScanThreads, B);
    if (stages & YS_STAGE_SCAN) {
        k_assign_scan<true><<<gs, kScanThreads, 0, stream>>>(L, io);
        YS_CHECK_LAUNCH("k_assign_scan");
    }
    if (stages & YS_STAGE_FIXUP) {
        k_assign_fixup<true><<<B, kFixThreads, 0, stream>>>(L, io, B);
        YS_CHECK_LAUNCH("k_assign_fixup");
    }

    LossParams P;
    P.tgi = io.tgi; P.norm = io.norm; P.gt_off = W.gt_off; P.gt_box = W.gt_box; P.gt_label = W.gt_label;
    P.tss = W.tss; P.partials = W.partials; P.nc = nc;
    P.hyp_box = hyp_box; P.hyp_cls = hyp_cls; P.hyp_dfl = hyp_dfl; P.batch = (float)B;
    dim3 gl(L.n_chunks, B);
    if (stages & YS_STAGE_LOSS) {
        if (level_grad_host)
            k_loss_fwd_bwd<true><<<gl, kLossThreads, 0, stream>>>(L, P);
        else
            k_loss_fwd_bwd<false><<<gl, kLossThreads, 0, stream>>>(L, P);
        YS_CHECK_LAUNCH("k_loss_fwd_bwd");
    }
    if (stages & YS_STAGE_FINAL) {
        k_loss_finalize<<<1, 1024, 0, stream>>>(W.partials, L.n_chunks * B, W.tss, hyp_box, hyp_cls, hyp_dfl, (float)B, loss_out);
        YS_CHECK_LAUNCH("k_loss_finalize");
    }
    return YS_OK;
}

extern "C" int ys_scale_inplace(float *data, size_t n, const float *scale_dev, void *stream_v) {
    YS_REQUIRE(data && scale_dev, YS_ERR_BAD_ARG, "null pointer");
    if (n == 0) return YS_OK;
    int blocks = (int)((n + 1023) / 1024);
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_scale_inplace<<<blocks, 256, 0, (cudaStream_t)stream_v>>>(data, n, scale_dev);
    YS_CHECK_LAUNCH("k_scale_inplace");
    return YS_OK;
}

// ---------------------------------------------------------------------------------------------------
// Dense assigner
// ---------------------------------------------------------------------------------------------------
namespace ys {
struct AssignWorkspace {
    int *owned_cnt, *gt_flags, *forced, *tgi;
    unsigned long long *best_key;
    float *norm, *tss;
    double *img_sum;
    unsigned *done_counter;
    size_t zero_bytes;   // leading region that must be zeroed per call
    size_t bytes;
};
static void carve_assign(AssignWorkspace &W, void *base, size_t cap, int B, int A, int G) {
    Carver c(base, cap);
    const size_t N = (size_t)B * G;
    W.done_counter = c.take<unsigned>(64);
    W.owned_cnt = c.take<int>(N);
    W.gt_flags = c.take<int>(N);
    W.best_key = c.take<unsigned long long>(N);
    W.zero_bytes = align_up(c.off, 256);
    W.tss = c.take<float>(64);
    W.forced = c.take<int>(2 * N);
    W.img_sum = c.take<double>((size_t)B);
    W.tgi = c.take<int>((size_t)B * A);
    W.norm = c.take<float>((size_t)B * A);
    W.bytes = align_up(c.off, 256);
}
}  // namespace ys

extern "C" size_t ys_assign_workspace_bytes(int B, int A, int G) {
    AssignWorkspace W;
    carve_assign(W, nullptr, 0, B > 0 ? B : 1, A > 0 ? A : 1, G > 0 ? G : 1);
    return W.bytes;
}

extern "C" int ys_assign(int B, int A, int G, int nc, const float *pd_scores, const float *pd_bboxes,
                         const float *anc_points, const float *gt_labels, const float *gt_bboxes,
                         const float *mask_gt, float alpha, float beta, float eps, float *target_labels,
                         float *target_bboxes, float *target_scores, uint8_t *fg_mask, int64_t *target_gt_idx,
                         float *norm_out, void *workspace, size_t workspace_bytes, void *stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    YS_REQUIRE(B >= 1 && A >= 1 && G >= 1 && nc >= 1, YS_ERR_BAD_ARG, "B, A, G, nc must be >= 1");
    YS_REQUIRE(B <= 65535, YS_ERR_UNSUPPORTED, "B must be <= 65535 (grid.y)");
    YS_REQUIRE(pd_scores && pd_bboxes && anc_points && gt_labels && gt_bboxes && mask_gt, YS_ERR_BAD_ARG, "null input");
    YS_REQUIRE(target_labels && target_bboxes && fg_mask && target_gt_idx && workspace, YS_ERR_BAD_ARG, "null output");
    YS_REQUIRE((((uintptr_t)pd_bboxes | (uintptr_t)gt_bboxes | (uintptr_t)target_bboxes) & 15) == 0 &&
                   ((uintptr_t)anc_points & 7) == 0, YS_ERR_BAD_ARG, "box pointers must be 16-byte aligned, anc_points 8-byte");
    AssignWorkspace W;
    carve_assign(W, workspace, workspace_bytes, B, A, G);
    YS_REQUIRE(W.bytes <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu", W.bytes, workspace_bytes);
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    cudaMemsetAsync(workspace, 0, W.zero_bytes, stream);

    LevelSet L;
    L.n = 1; L.A = A; L.C = nc; L.n_chunks = 0;
    for (int i = 0; i < YS_MAX_LEVELS; ++i) { L.x[i] = nullptr; L.g[i] = nullptr; L.w[i] = 1; L.hw[i] = 0; L.stride[i] = 1.f; }
    for (int i = 0; i <= YS_MAX_LEVELS; ++i) { L.a_off[i] = i == 0 ? 0 : A; L.chunk_off[i] = 0; }

    AssignIO io;
    io.gt_off = nullptr; io.G = G; io.gt_box = (const float4 *)gt_bboxes; io.gt_label_i = nullptr;
    io.gt_label_f = gt_labels; io.gt_valid = mask_gt; io.pd_box = (const float4 *)pd_bboxes; io.pd_scores = pd_scores;
    io.anc = (const float2 *)anc_points; io.best_key = W.best_key; io.owned_cnt = W.owned_cnt; io.gt_flags = W.gt_flags;
    io.forced = W.forced; io.pbox = nullptr; io.tgi = W.tgi; io.norm = norm_out ? norm_out : W.norm;
    io.img_sum = W.img_sum; io.tss = W.tss; io.done_counter = W.done_counter;
    io.nc = nc; io.alpha = alpha; io.beta = beta; io.eps = eps;

    dim3 gs((A + kScanThreads - 1) / kScanThreads, B);
    k_assign_scan<false><<<gs, kScanThreads, 0, stream>>>(L, io);
    YS_CHECK_LAUNCH("k_assign_scan(dense)");
    k_assign_fixup<false><<<B, kFixThreads, 0, stream>>>(L, io, B);
    YS_CHECK_LAUNCH("k_assign_fixup(dense)");
    const size_t tot = (size_t)B * A;
    k_assign_emit<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(B, A, G, nc, W.tgi, io.norm, gt_labels,
                                                                   (const float4 *)gt_bboxes, target_labels,
                                                                   (float4 *)target_bboxes, target_scores, fg_mask,
                                                                   (long long *)target_gt_idx);
    YS_CHECK_LAUNCH("k_assign_emit");
    return YS_OK;
}
