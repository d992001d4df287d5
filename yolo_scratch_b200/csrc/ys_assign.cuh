// ys_assign.cuh -- task-aligned assignment kernels, shared by the fused loss
// (boxes decoded on the fly from the NCHW head levels, GT as a CSR list) and by the dense
// TaskAlignedAssigner.forward entry point (boxes / scores / padded GT given).
//
// Replaces losses/yolo_v8_loss.py:78-270 of the reference.  Structure:
//   k_assign_scan   one thread per anchor: decode (fused), loop over the image's GT staged in shared
//                   memory; per anchor: in-box count, first in-box GT, first-max IoU GT -> preliminary
//                   assignment; per GT (cross-anchor facts the reference gets from its B x G Python
//                   loops): any anchor inside, owned-anchor count, and first-max-IoU anchor as an
//                   order-preserving u64 atomicMax key.
//   k_assign_fixup  one block per image: the forced assignments of the reference's two loops
//                   (:117-139, :146-162) replayed with exact sequential semantics from the per-GT facts,
//                   touching only the handful of forced anchors; then the per-image sum of target scores
//                   and (last block) target_scores_sum.
#pragma once
#include "ys_common.cuh"

namespace ys {

constexpr int kScanThreads = 256;
constexpr int kGtTile = 256;        // GT boxes staged per shared-memory tile
constexpr int kFixThreads = 256;
constexpr int kFixSmemGt = 2048;    // per-image GT count handled with shared-memory counters in the fix-up

// Everything the assignment kernels need to know about where boxes, scores and GT come from.
struct AssignIO {
    // GT (CSR for fused: off = gt_off[b]; dense: off = b*G, count = G)
    const int *gt_off;          // fused only
    int G;                      // dense only
    const float4 *gt_box;       // px xyxy
    const int *gt_label_i;      // fused: int labels
    const float *gt_label_f;    // dense: float labels
    const float *gt_valid;      // dense: mask_gt (B,G); fused: nullptr (all valid)
    // predictions (dense only)
    const float4 *pd_box;       // (B,A) px xyxy
    const float *pd_scores;     // (B,A,nc) probabilities
    const float2 *anc;          // (A) px
    // per-GT scratch, indexed like gt_box
    unsigned long long *best_key;
    int *owned_cnt;
    int *gt_flags;              // bit0: some anchor centre lies inside
    int *forced;                // 2 ints per GT: anchor forced by loop 1 / loop 2, -1 = none
    // per-anchor outputs (B,A)
    float4 *pbox;               // fused: decoded px boxes (written by the scan)
    int *tgi;                   // -1 = background
    float *norm;
    // reductions
    double *img_sum;            // (B)
    float *tss;                 // [0] = max(sum norm, 1)
    unsigned *done_counter;
    int nc;
    float alpha, beta, eps;
};

__device__ __forceinline__ int level_of(const LevelSet &L, int a) {
    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && a >= L.a_off[k]) l = k;
    return l;
}

// Softmax expectation over 16 logits: d = sum_j j*softmax(z)_j  (losses/yolo_v8_loss.py:534-536).
// Also returns the max and the sum of exponentials for the DFL cross-entropy.
__device__ __forceinline__ float dfl_expect16(const float (&z)[16], float &zmax, float &esum) {
    float m = z[0];
#pragma unroll
    for (int j = 1; j < 16; ++j) m = fmaxf(m, z[j]);
    float s = 0.f, d = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        float e = expf(z[j] - m);
        s += e;
        d += e * (float)j;
    }
    zmax = m;
    esum = s;
    return d / s;
}

template <bool FUSED>
__device__ __forceinline__ void gt_range(const AssignIO &io, int b, int &g0, int &n) {
    if (FUSED) {
        g0 = io.gt_off[b];
        n = io.gt_off[b + 1] - g0;
    } else {
        g0 = b * io.G;
        n = io.G;
    }
}

// Score of anchor a for class `label` (probability).  Fused: sigmoid of the class logit in the level tensor.
template <bool FUSED>
__device__ __forceinline__ float score_at(const AssignIO &io, const LevelSet &L, int b, int a, int label) {
    if (FUSED) {
        int l = level_of(L, a);
        int i = a - L.a_off[l];
        int hw = L.hw[l];
        float x = __ldg(L.x[l] + ((size_t)b * L.C + (L.C - io.nc) + label) * hw + i);
        return sigmoidf(x);
    } else {
        return __ldg(io.pd_scores + ((size_t)b * L.A + a) * io.nc + label);
    }
}

template <bool FUSED>
__device__ __forceinline__ int label_of(const AssignIO &io, int gidx) {
    int lab = FUSED ? io.gt_label_i[gidx] : (int)io.gt_label_f[gidx];
    return lab;
}

template <bool FUSED>
__global__ void __launch_bounds__(kScanThreads) k_assign_scan(LevelSet L, AssignIO io) {
    __shared__ float4 s_box[kGtTile];
    __shared__ float s_area[kGtTile];
    __shared__ unsigned long long s_best[kGtTile];
    __shared__ int s_any[kGtTile];
    __shared__ unsigned char s_valid[kGtTile];

    const int b = blockIdx.y;
    const int a = blockIdx.x * kScanThreads + threadIdx.x;
    const bool live = a < L.A;

    int g0, n;
    gt_range<FUSED>(io, b, g0, n);

    // ---- this anchor's predicted box (px) and centre (px)
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    float ax = 0.f, ay = 0.f;
    if (live) {
        if (FUSED) {
            int l = level_of(L, a);
            int i = a - L.a_off[l];
            int w = L.w[l], hw = L.hw[l];
            float st = L.stride[l];
            int iy = i / w, ix = i - iy * w;
            float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
            const float *base = L.x[l] + (size_t)b * L.C * hw + i;
            float d[4];
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                float z[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) z[j] = __ldg(base + (size_t)(s * 16 + j) * hw);
                float m, es;
                d[s] = dfl_expect16(z, m, es);
            }
            // dist2bbox in grid units, then * stride (losses/yolo_v8_loss.py:558-559, :412)
            p.x = YS_MUL(YS_SUB(gx, d[0]), st);
            p.y = YS_MUL(YS_SUB(gy, d[1]), st);
            p.z = YS_MUL(YS_ADD(gx, d[2]), st);
            p.w = YS_MUL(YS_ADD(gy, d[3]), st);
            ax = YS_MUL(gx, st);
            ay = YS_MUL(gy, st);
            io.pbox[(size_t)b * L.A + a] = p;
        } else {
            p = io.pd_box[(size_t)b * L.A + a];
            float2 c = io.anc[a];
            ax = c.x;
            ay = c.y;
        }
    }
    const float area_p = area_eps_h(p.x, p.y, p.z, p.w);

    int cnt = 0, first = -1, best_g = 0;
    float best_ov = -1.f, first_ov = 0.f;

    for (int t0 = 0; t0 < n; t0 += kGtTile) {
        const int tn = min(kGtTile, n - t0);
        __syncthreads();
        for (int j = threadIdx.x; j < tn; j += kScanThreads) {
            float4 q = io.gt_box[g0 + t0 + j];
            s_box[j] = q;
            s_area[j] = area_eps_h(q.x, q.y, q.z, q.w);
            s_best[j] = 0ull;
            s_any[j] = 0;
            s_valid[j] = FUSED ? 1 : (io.gt_valid[g0 + t0 + j] != 0.f);
        }
        __syncthreads();
        if (live) {
            for (int j = 0; j < tn; ++j) {
                const float4 q = s_box[j];
                const float ov = iou_assign(p, area_p, q, s_area[j]);
                const bool in = centre_in_box(ax, ay, q) && s_valid[j];
                if (in) {
                    if (first < 0) { first = t0 + j; first_ov = ov; }
                    ++cnt;
                    s_any[j] = 1;
                }
                if (ov > best_ov) { best_ov = ov; best_g = t0 + j; }
                const unsigned long long key = argmax_key(ov, (unsigned)a);
                if (ov >= 0.f && key > s_best[j]) atomicMax(&s_best[j], key);
            }
        }
        __syncthreads();
        for (int j = threadIdx.x; j < tn; j += kScanThreads) {
            if (s_best[j]) atomicMax(&io.best_key[g0 + t0 + j], s_best[j]);
            if (s_any[j]) io.gt_flags[g0 + t0 + j] = 1;
        }
    }

    if (!live) return;
    // ---- preliminary per-anchor resolve (no forced entries yet): losses/yolo_v8_loss.py:226-244
    int g = -1;
    float ov = 0.f;
    if (cnt > 1) { g = best_g; ov = best_ov; }
    else if (cnt == 1) { g = first; ov = first_ov; }
    float nrm = 0.f;
    if (g >= 0) {
        atomicAdd(&io.owned_cnt[g0 + g], 1);
        float sc = score_at<FUSED>(io, L, b, a, label_of<FUSED>(io, g0 + g));
        nrm = align_norm(sc, ov, io.alpha, io.beta, io.eps);
    }
    io.tgi[(size_t)b * L.A + a] = g;
    io.norm[(size_t)b * L.A + a] = nrm;
}

// Recompute one anchor's row of mask_pos / overlaps with the forced entries of loops 1 (and 2),
// warp-parallel over the image's GT.  Returns the resolved GT (>= 0) and its overlap.
template <bool FUSED>
__device__ __forceinline__ int resolve_row(const AssignIO &io, const LevelSet &L, int b, int k, int g0, int n,
                                           const int *forced, bool with_loop2, float &ov_out) {
    const int lane = threadIdx.x & 31;
    float4 p;
    float ax, ay;
    if (FUSED) {
        p = io.pbox[(size_t)b * L.A + k];
        int l = level_of(L, k);
        int i = k - L.a_off[l];
        int iy = i / L.w[l], ix = i - iy * L.w[l];
        ax = YS_MUL((float)ix + 0.5f, L.stride[l]);
        ay = YS_MUL((float)iy + 0.5f, L.stride[l]);
    } else {
        p = io.pd_box[(size_t)b * L.A + k];
        float2 c = io.anc[k];
        ax = c.x;
        ay = c.y;
    }
    const float area_p = area_eps_h(p.x, p.y, p.z, p.w);
    int cnt = 0;
    unsigned first = 0xFFFFFFFFu;
    unsigned long long best = 0ull;   // argmax over ALL slots of ov, first max
    for (int j = lane; j < n; j += 32) {
        float4 q = io.gt_box[g0 + j];
        float ov = iou_assign(p, area_p, q, area_eps_h(q.x, q.y, q.z, q.w));
        bool valid = FUSED ? true : (io.gt_valid[g0 + j] != 0.f);
        bool m = (centre_in_box(ax, ay, q) && valid) || forced[2 * j] == k || (with_loop2 && forced[2 * j + 1] == k);
        if (m) { ++cnt; first = min(first, (unsigned)j); }
        unsigned long long key = argmax_key(ov, (unsigned)j);
        if (ov >= 0.f && key > best) best = key;
    }
    cnt = warp_sum(cnt);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    best = warp_max(best);
    int g = -1;
    if (cnt > 1) g = (int)argmax_key_index(best);
    else if (cnt == 1) g = (int)first;
    float ov = 0.f;
    if (g >= 0) {
        float4 q = io.gt_box[g0 + g];
        ov = iou_assign(p, area_p, q, area_eps_h(q.x, q.y, q.z, q.w));
    }
    ov_out = ov;
    return g;
}

template <bool FUSED>
__global__ void __launch_bounds__(kFixThreads) k_assign_fixup(LevelSet L, AssignIO io, int B) {
    __shared__ int s_cnt[kFixSmemGt];
    __shared__ int s_forced[2 * kFixSmemGt];
    __shared__ int s_todo;
    __shared__ double s_red[kFixThreads / 32];
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int g0, n;
    gt_range<FUSED>(io, b, g0, n);
    int *tgi = io.tgi + (size_t)b * L.A;
    float *norm = io.norm + (size_t)b * L.A;
    // per-GT scratch: shared memory when the image's GT fit, else the global scratch in place
    const bool in_smem = n <= kFixSmemGt;
    int *cntp = in_smem ? s_cnt : (io.owned_cnt + g0);
    int *forced = in_smem ? s_forced : (io.forced + 2 * (size_t)g0);   // [2j] loop-1 anchor, [2j+1] loop-2 anchor

    // ---- does this image need any forced assignment?  (common case: no)
    if (tid == 0) s_todo = 0;
    __syncthreads();
    for (int j = tid; j < n; j += kFixThreads) {
        bool valid = FUSED ? true : (io.gt_valid[g0 + j] != 0.f);
        int c = io.owned_cnt[g0 + j];
        if (in_smem) s_cnt[j] = c;
        bool lonely1 = valid && io.gt_flags[g0 + j] == 0;
        // loop 1 (losses/yolo_v8_loss.py:117-139): valid GT with no anchor inside -> its best-IoU anchor
        forced[2 * j] = lonely1 ? (int)argmax_key_index(io.best_key[g0 + j]) : -1;
        forced[2 * j + 1] = -1;
        if (valid && (lonely1 || c == 0)) s_todo = 1;
    }
    __syncthreads();
    if (s_todo && warp == 0) {
        // -- first resolve (:142) for the anchors loop 1 touched; everyone else keeps the scan's result
        for (int j = 0; j < n; ++j) {
            int k = forced[2 * j];
            if (k < 0) continue;
            bool seen = false;   // k already handled through an earlier GT?
            for (int jj = lane; jj < j; jj += 32) seen |= (forced[2 * jj] == k);
            if (__any_sync(0xffffffffu, seen)) continue;
            float ov;
            int gnew = resolve_row<FUSED>(io, L, b, k, g0, n, forced, false, ov);
            if (lane == 0) {
                int gold = tgi[k];
                if (gold >= 0) cntp[gold] -= 1;
                if (gnew >= 0) cntp[gnew] += 1;
                tgi[k] = gnew;
            }
            __syncwarp();
        }
        // -- loop 2 (:146-162): sequential in g; later GT see the edits of earlier ones
        if (lane == 0) {
            for (int j = 0; j < n; ++j) {
                bool valid = FUSED ? true : (io.gt_valid[g0 + j] != 0.f);
                if (!valid || cntp[j] != 0) continue;
                int k = (int)argmax_key_index(io.best_key[g0 + j]);
                forced[2 * j + 1] = k;
                int gold = tgi[k];
                if (gold >= 0) cntp[gold] -= 1;
                tgi[k] = j;
                cntp[j] += 1;
            }
        }
        __syncwarp();
        // -- final resolve (:165) from mask_pos alone, for every anchor that carries a forced entry
        for (int j = 0; j < 2 * n; ++j) {
            int k = forced[j];
            if (k < 0) continue;
            bool seen = false;
            for (int jj = lane; jj < j; jj += 32) seen |= (forced[jj] == k);
            if (__any_sync(0xffffffffu, seen)) continue;
            float ov;
            int gfin = resolve_row<FUSED>(io, L, b, k, g0, n, forced, true, ov);
            if (lane == 0) {
                tgi[k] = gfin;
                float nrm = 0.f;
                if (gfin >= 0) {
                    float sc = score_at<FUSED>(io, L, b, k, label_of<FUSED>(io, g0 + gfin));
                    nrm = align_norm(sc, ov, io.alpha, io.beta, io.eps);
                }
                norm[k] = nrm;
            }
            __syncwarp();
        }
    }
    __syncthreads();

    // ---- per-image sum of target scores (fixed order), then target_scores_sum in the last block
    double acc = 0.0;
    for (int a = tid; a < L.A; a += kFixThreads) acc += (double)norm[a];
    acc = warp_sum(acc);
    if (lane == 0) s_red[warp] = acc;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < kFixThreads / 32; ++w) t += s_red[w];
        io.img_sum[b] = t;
        __threadfence();
        unsigned prev = atomicAdd(io.done_counter, 1u);
        if (prev == (unsigned)B - 1u) {
            __threadfence();
            double tot = 0.0;
            for (int i = 0; i < B; ++i) tot += ((volatile double *)io.img_sum)[i];
            float tss = (float)tot;
            io.tss[0] = tss > 1.f ? tss : 1.f;     // max(target_scores.sum(), 1)  (:472)
            *io.done_counter = 0u;
        }
    }
}

}  // namespace ys
