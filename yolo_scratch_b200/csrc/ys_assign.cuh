// ys_assign.cuh -- task-aligned assignment, shared by the fused loss (boxes decoded on the
// fly from the NCHW head levels, GT as a CSR list) and by the dense TaskAlignedAssigner.forward entry point
// (boxes / scores / padded GT given).  Replaces losses/yolo_v8_loss.py:78-270 of the reference.
//
// ONE kernel, k_assign:
//   scan phase   one thread per anchor; the image's GT staged in shared memory.  Pass 0 is the cheap one:
//                centre-in-box test per GT (count, first hit, per-GT "someone is inside" flag).  Only anchors
//                that sit inside more than one GT run pass 1, the IoU arg-max over all GT slots (:226-244);
//                the IEEE division is executed only for pairs whose boxes intersect.  Emits the preliminary
//                assignment, its target score (norm), per-GT owner counts and a per-block partial sum of norm.
//   fix-up phase run by the LAST block of each image to finish (threadfence + per-image arrival counter):
//                the forced assignments of the reference's two Python loops (:117-139, :146-162) replayed with
//                their exact sequential semantics, touching only the handful of forced anchors.  The per-GT
//                "best-IoU anchor" (arg-max over all anchors) is computed here, lazily, only for the GT that
//                need it.  Then the image's sum of target scores; the last image computes target_scores_sum.
#pragma once
#include <cuda.h>
#include "ys_common.cuh"

namespace ys {

#ifndef YS_SCAN_THREADS
#define YS_SCAN_THREADS 256
#endif
#ifndef YS_SCAN_MINBLOCKS
#define YS_SCAN_MINBLOCKS 4
#endif
constexpr int kScanThreads = YS_SCAN_THREADS;
#ifndef YS_GT_TILE
#define YS_GT_TILE (YS_SCAN_THREADS < 256 ? YS_SCAN_THREADS : 256)
#endif
constexpr int kGtTile = YS_GT_TILE;        // GT boxes staged per shared-memory tile
constexpr int kFixSmemGt = 256;     // per-image GT count the fix-up handles with shared-memory scratch (dense entry point)
#ifndef YS_SCAN_STAGES
#define YS_SCAN_STAGES 2
#endif
constexpr int kScanStages = YS_SCAN_STAGES;       // fused path: per-warp ring of side tiles (16 DFL bins x 32 anchors, 2 KB each)
constexpr int kScanSideBytes = 16 * 32 * 4;
constexpr int kScanTileBytes = (kScanThreads / 32) * kScanStages * kScanSideBytes;   // all warps' rings: 32 KB
constexpr int kFixSmemGtFused = (kScanTileBytes / 36) & ~63;   // per-image GT count of the fused fix-up's shared-memory
                                                               // scratch: it lives in the (consumed) rings, 36 B per GT
constexpr float kDflReach = 15.5f;  // a decoded side is an expectation over bins 0..15: a predicted box reaches at most
                                    // 15 cells from its anchor centre (0.5 cell of slack for the approximate softmax)

// Fused path: one 2-D tensor map per level over [B*C rows][h*w columns] fp32, box {32 columns, 16 rows}: one side
// (16 DFL bins) of one warp's 32 anchors per TMA operation.
struct ScanMaps {
    CUtensorMap lvl[YS_MAX_LEVELS];
};

// Everything the assignment kernel needs to know about where boxes, scores and GT come from.
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
    // per-GT scratch, indexed like gt_box (zero on entry; re-zeroed by the fix-up)
    int *owned_cnt;
    int *gt_flags;              // 1: some anchor centre lies inside
    const int *img_order;       // fused only (else nullptr): blockIdx.y -> image, a scheduling hint from k_pack_gt
    int n_identity;             // fused only: the first n_identity launch slots are the images 0..n_identity-1 (see k_assign)
    int skip_fixup;             // profiling only
    // dense entry point only, non-default modes (north_star wording; the reference runs neither)
    int ciou;                   // 1: pair metric = clamp(CIoU, 0) (reference bbox_iou(CIoU=True)) instead of plain IoU
    const unsigned *sel;        // top-k mode: bit a of row (b*G + g): anchor a is among GT g's top-k candidates; nullptr = off
    int sel_words;              // words per row = ceil(A / 32)
    unsigned long long *best_key;   // per-GT arg-max key over all anchors: written by the scan for risk GT (zero on
                                    // entry), by the fix-up for the others when an image has more than kFixSmemGt GT
    int *forced;                // same: 8 ints per GT (forced x2, needy, kk, tg0, lost, spare x2)
    // per-anchor outputs (B,A)
    float4 *pbox;               // fused: decoded px boxes (written by the scan)
    int *tgi;                   // -1 = background
    int *lab;                   // class of the assigned GT, -1 = background (may be nullptr)
    float *norm;
    // reductions
    float *partial;             // (B, gridDim.x) per-block sums of norm
    double *img_sum;            // (B)
    float *tss;                 // [0] = max(sum norm, 1)
    unsigned *img_done;         // (B) arrival counters, self-resetting
    unsigned *done_counter;     // images finished, self-resetting
    int nc;
    float alpha, beta, eps;
};

__device__ __forceinline__ unsigned float_to_ord(float f) {
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord_to_float(unsigned o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__device__ __forceinline__ int level_of(const LevelSet &L, int a) {
    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && a >= L.a_off[k]) l = k;
    return l;
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
        return sigmoid_fast(x);
    } else {
        return __ldg(io.pd_scores + ((size_t)b * L.A + a) * io.nc + label);
    }
}

template <bool FUSED>
__device__ __forceinline__ int label_of(const AssignIO &io, int gidx) {
    return FUSED ? io.gt_label_i[gidx] : (int)io.gt_label_f[gidx];
}

// Assigner overlap with the division skipped when the boxes do not intersect: inter == +0 then, the union is
// positive for well-formed boxes (x2 >= x1, y2 >= y1) and 0/union == +0 bit-for-bit.
__device__ __forceinline__ float iou_assign_lazy(float4 p, float area_p, float4 q, float area_q) {
    float iw = YS_SUB(fminf2(p.z, q.z), fmaxf2(p.x, q.x));
    float ih = YS_SUB(fminf2(p.w, q.w), fmaxf2(p.y, q.y));
    if (!(iw > 0.f) || !(ih > 0.f)) return 0.f;
    const float inter = YS_MUL(iw, ih);
    const float uni = YS_ADD(YS_SUB(YS_ADD(area_p, area_q), inter), kEpsIou);
    const float iou = YS_DIV(inter, uni);
    return iou < 0.f ? 0.f : iou;
}

// The assigner's pair metric.  Fused path: always the reference's plain IoU (:199).  Dense entry point: clamp(CIoU, 0) in
// ciou mode -- CIoU <= IoU, so it clamps to 0 wherever the boxes do not intersect, exactly like the lazy IoU.
template <bool FUSED, bool LAZY = true>
__device__ __forceinline__ float pair_ov(const AssignIO &io, float4 p, float area_p, float4 q, float area_q) {
    if (!FUSED && io.ciou) {
        const float iw = YS_SUB(fminf2(p.z, q.z), fmaxf2(p.x, q.x));
        const float ih = YS_SUB(fminf2(p.w, q.w), fmaxf2(p.y, q.y));
        if (!(iw > 0.f) || !(ih > 0.f)) return 0.f;
        const float c = ciou_pair(p, q);
        return c < 0.f ? 0.f : c;
    }
    return LAZY ? iou_assign_lazy(p, area_p, q, area_q) : iou_assign(p, area_p, q, area_q);
}
// Top-k mode (dense entry point): is anchor a among the selected candidates of GT slot j of image b?
template <bool FUSED>
__device__ __forceinline__ bool sel_ok(const AssignIO &io, int gidx, int a) {
    if (FUSED || io.sel == nullptr) return true;
    return (io.sel[(size_t)gidx * io.sel_words + (a >> 5)] >> (a & 31)) & 1u;
}

// This anchor's centre (px) and, for the fused path, its grid position.
template <bool FUSED>
__device__ __forceinline__ void anchor_centre(const AssignIO &io, const LevelSet &L, int a, float &ax, float &ay) {
    if (FUSED) {
        int l = level_of(L, a);
        int i = a - L.a_off[l];
        int iy = i / L.w[l], ix = i - iy * L.w[l];
        ax = YS_MUL((float)ix + 0.5f, L.stride[l]);
        ay = YS_MUL((float)iy + 0.5f, L.stride[l]);
    } else {
        float2 c = io.anc[a];
        ax = c.x;
        ay = c.y;
    }
}

// Recompute one anchor's row of mask_pos / overlaps with the forced entries of loops 1 (and 2),
// warp-parallel over the image's GT.  Returns the resolved GT (>= 0, or -1) and its overlap.
template <bool FUSED>
__device__ __forceinline__ int resolve_row(const AssignIO &io, const LevelSet &L, int b, int k, int g0, int n,
                                           const int *forced, bool with_loop2, float &ov_out) {
    const int lane = threadIdx.x & 31;
    const float4 p = FUSED ? __ldcg(io.pbox + (size_t)b * L.A + k) : io.pd_box[(size_t)b * L.A + k];
    float ax, ay;
    anchor_centre<FUSED>(io, L, k, ax, ay);
    const float area_p = area_eps_h(p.x, p.y, p.z, p.w);
    int cnt = 0;
    unsigned first = 0xFFFFFFFFu;
    unsigned long long best = 0ull;   // argmax over ALL slots of ov, first max
    for (int j = lane; j < n; j += 32) {
        float4 q = io.gt_box[g0 + j];
        float ov = pair_ov<FUSED, false>(io, p, area_p, q, area_eps_h(q.x, q.y, q.z, q.w));
        bool valid = FUSED ? true : (io.gt_valid[g0 + j] != 0.f);
        bool m = (centre_in_box(ax, ay, q) && valid && sel_ok<FUSED>(io, g0 + j, k)) || forced[2 * j] == k || (with_loop2 && forced[2 * j + 1] == k);
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
        ov = pair_ov<FUSED, false>(io, p, area_p, q, area_eps_h(q.x, q.y, q.z, q.w));
    }
    ov_out = ov;
    return g;
}

// The same row resolve (loops 1 and 2 included) by a group of W lanes (W a power of two, 1..32; groups are aligned
// lane ranges of a warp, every lane of the warp calls this).  The final resolve picks W from the number of rows it
// has to redo: few rows -> a warp each, hundreds (dense scenes) -> a thread each, so the GT loop is never the
// serial part.  Overlaps go through the lazy form (no division for disjoint boxes, same bits).
template <bool FUSED>
__device__ __forceinline__ int resolve_row_group(const AssignIO &io, const LevelSet &L, int b, int k, int g0, int n,
                                                 const int *forced, int W, bool active, float &ov_out) {
    const int sub = threadIdx.x & (W - 1);
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    float ax = 0.f, ay = 0.f, area_p = 0.f;
    int cnt = 0;
    unsigned first = 0xFFFFFFFFu;
    unsigned long long best = 0ull;
    if (active) {
        p = FUSED ? __ldcg(io.pbox + (size_t)b * L.A + k) : io.pd_box[(size_t)b * L.A + k];
        anchor_centre<FUSED>(io, L, k, ax, ay);
        area_p = area_eps_h(p.x, p.y, p.z, p.w);
        float best_ov = -1.f;      // this lane's first maximum (ascending j)
        int best_j = 0;
        // four GT per trip, their loads issued together (the loop is latency-bound: one warp, dependent loads)
        for (int j0 = sub; j0 < n; j0 += 4 * W) {
            float4 q[4];
            int2 f[4];
            bool valid[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int jc = min(j0 + u * W, n - 1);
                q[u] = io.gt_box[g0 + jc];
                f[u] = *reinterpret_cast<const int2 *>(forced + 2 * jc);
                valid[u] = FUSED ? true : (io.gt_valid[g0 + jc] != 0.f);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u * W;
                if (j < n) {
                    const float ov = pair_ov<FUSED>(io, p, area_p, q[u], area_eps_h(q[u].x, q[u].y, q[u].z, q[u].w));
                    const bool m = (centre_in_box(ax, ay, q[u]) && valid[u] && sel_ok<FUSED>(io, g0 + j, k)) || f[u].x == k || f[u].y == k;
                    if (m) { ++cnt; first = min(first, (unsigned)j); }
                    if (ov > best_ov) { best_ov = ov; best_j = j; }
                }
            }
        }
        if (best_ov >= 0.f) best = argmax_key(best_ov, (unsigned)best_j);
    }
    for (int o = W >> 1; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other > best ? other : best;
    }
    int g = -1;
    if (cnt > 1) g = (int)argmax_key_index(best);
    else if (cnt == 1) g = (int)first;
    float ov = 0.f;
    if (active && g >= 0) {
        const float4 q = io.gt_box[g0 + g];
        ov = pair_ov<FUSED, false>(io, p, area_p, q, area_eps_h(q.x, q.y, q.z, q.w));
    }
    ov_out = ov;
    return g;
}

// Shared-memory layout: the scan phase and the fix-up phase never overlap in time, so they alias one buffer.
struct ScanSmem {
    float4 box[kGtTile];
    float area[kGtTile];
    int any[kGtTile];
    int own[kGtTile];
    unsigned long long rkey[kGtTile];   // fused: this block's arg-max key (best IoU, lowest anchor) per GT of the tile; 0 = none
    unsigned char valid[kGtTile];
};
struct FixSmem {
    unsigned long long key[kFixSmemGt];
    int cnt[kFixSmemGt];
    int forced[2 * kFixSmemGt];
    int needy[kFixSmemGt];
    int kk[kFixSmemGt];
    int tg0[kFixSmemGt];
    int lost[kFixSmemGt];
};
union AssignSmem {
    ScanSmem s;
    FixSmem f;
};

template <bool FUSED>
__device__ void fixup_image(const LevelSet &L, const AssignIO &io, int b, int B, void *scratch, int cap, int nblk);

// One side tile (16 DFL bins x the warp's 32 anchors) into stage `st` of the warp's ring; the stage's mbarrier gets the bytes.
__device__ __forceinline__ void scan_issue_side(unsigned char *s_dyn, unsigned long long *mbar, const CUtensorMap *map, int warp,
                                                int st, int side, int i0, int row0) {
    const unsigned bar = (unsigned)__cvta_generic_to_shared(mbar);
    const unsigned dst = (unsigned)__cvta_generic_to_shared(s_dyn + (size_t)(warp * kScanStages + st) * kScanSideBytes);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kScanSideBytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(i0 + 32 * warp), "r"(row0 + 16 * side), "r"(bar) : "memory");
}

#ifdef YS_SCANTIME        // timing experiment: thread 0 of a sample of blocks prints the block's phase stamps (ns)
#define YS_BSTAMP(i) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); bst[i] = t_; } } while (0)
#else
#define YS_BSTAMP(i) do { } while (0)
#endif
template <bool FUSED>
__global__ void __launch_bounds__(kScanThreads, YS_SCAN_MINBLOCKS)
k_assign(LevelSet L, AssignIO io, int B, const __grid_constant__ ScanMaps M) {
    // dynamic shared memory: [fused: per-warp rings of side tiles, kScanStages x 2 KB each][AssignSmem]
    extern __shared__ __align__(128) unsigned char s_dyn[];
    AssignSmem &sm = *reinterpret_cast<AssignSmem *>(s_dyn + (FUSED ? kScanTileBytes : 0));
    __shared__ __align__(8) unsigned long long s_mbar[kScanThreads / 32][kScanStages];
    __shared__ float s_red[kScanThreads / 32];
    __shared__ int s_flag;
    ScanSmem &S = sm.s;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef YS_SCANTIME
    unsigned long long bst[8] = {};
#endif
    YS_BSTAMP(0);
    // Fused path, programmatic dependent launch: this grid becomes resident while k_pack_gt still runs.  The blocks of
    // the first io.n_identity images (launch order == image index for them, by agreement with k_pack_gt: about one
    // wave of blocks) do not need anything k_pack_gt writes until after the decode, so they fetch and decode their
    // tile first and wait for k_pack_gt then; every other block waits here (it needs the image order).
    const bool early = FUSED && (int)blockIdx.y < io.n_identity;
    if (FUSED) {
        cudaTriggerProgrammaticLaunchCompletion();      // k_loss_stream may become resident behind this grid's last wave
        if (!early) cudaGridDependencySynchronize();
#ifdef YS_FIXTIME
        if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
            unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
            printf("START t=%llu\n", t_);
        }
#endif
    }
    YS_BSTAMP(1);
    const int b = (FUSED && io.img_order && !early) ? io.img_order[blockIdx.y] : (int)blockIdx.y;
    // fused: blocks are cut per level (a block never straddles two levels: one tensor operation loads its tile)
    int lvl = 0, i_lvl = 0, a;
    bool live, warp_cols = false, use_tma = false;
    if (FUSED) {
#pragma unroll
        for (int k = 1; k < YS_MAX_LEVELS; ++k)
            if (k < L.n && (int)blockIdx.x >= L.sblk_off[k]) lvl = k;
        const int i0 = ((int)blockIdx.x - L.sblk_off[lvl]) * kScanThreads;
        i_lvl = i0 + tid;
        live = i_lvl < L.hw[lvl];
        a = L.a_off[lvl] + i_lvl;
        // each warp feeds itself: lane 0 starts the first kScanStages sides of the warp's 32 anchors
        warp_cols = i0 + 32 * warp < L.hw[lvl];
        use_tma = L.tma[lvl] != 0;      // rows of this level 16-byte aligned: side tiles by TMA; else plain loads below
        if (lane == 0 && warp_cols && use_tma) {
#pragma unroll
            for (int s = 0; s < kScanStages; ++s) {
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_mbar[warp][s]);
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#pragma unroll
            for (int s = 0; s < kScanStages && s < 4; ++s) scan_issue_side(s_dyn, &s_mbar[warp][s], &M.lvl[lvl], warp, s, s, i0, b * L.C);
        }
        __syncwarp();
    } else {
        a = blockIdx.x * kScanThreads + tid;
        live = a < L.A;
    }

    int g0 = 0, n = 0;
    if (!early) gt_range<FUSED>(io, b, g0, n);
    int ntiles = (n + kGtTile - 1) / kGtTile;
    bool one_tile = ntiles <= 1;
    // GT tile -> shared memory (no barriers inside; the caller brackets it)
    auto stage_gt = [&](int t0, int tn) {
        for (int j = tid; j < tn; j += kScanThreads) {
            const float4 q = io.gt_box[g0 + t0 + j];
            S.box[j] = q;
            S.area[j] = area_eps_h(q.x, q.y, q.z, q.w);
            S.any[j] = 0;
            S.own[j] = 0;
            S.rkey[j] = 0ull;
            S.valid[j] = FUSED ? 1 : (io.gt_valid[g0 + t0 + j] != 0.f);
        }
    };
    __syncthreads();               // (fused) mbarrier initialised
    // GT tile 0: one GT per thread goes into a register now and into shared memory after the decode below, so the
    // latency of this (dependent) load hides under the decode instead of in front of a barrier
    static_assert(kGtTile <= kScanThreads, "tile 0 is staged one GT per thread");
    int tn0 = ntiles > 0 ? min(kGtTile, n) : 0;
    float4 q_pre = make_float4(0.f, 0.f, 0.f, 0.f);
    bool v_pre = true;
    if (tid < tn0) {
        q_pre = io.gt_box[g0 + tid];
        if (!FUSED) v_pre = io.gt_valid[g0 + tid] != 0.f;
    }

    // ---- this anchor's predicted box (px) and centre (px)
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    float ax = 0.f, ay = 0.f;
    if (FUSED) {
        float d[4] = {0.f, 0.f, 0.f, 0.f};
        if (warp_cols && !use_tma) {
            // plain-load loader (h*w % 4 != 0 or unaligned base: no tensor map over this level): lane = anchor, every
            // channel row piece of the warp is one contiguous 128-byte run
            const float *xs = L.x[lvl] + (size_t)b * L.C * L.hw[lvl] + i_lvl;
            const size_t hw = (size_t)L.hw[lvl];
#pragma unroll 1
            for (int s = 0; s < 4; ++s) {
                float z[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) z[j] = live ? __ldg(xs + (size_t)(16 * s + j) * hw) : 0.f;
                float m2, inv_s;
                const float v = dfl_stats16(z, m2, inv_s);
                d[0] = s == 0 ? v : d[0]; d[1] = s == 1 ? v : d[1]; d[2] = s == 2 ? v : d[2]; d[3] = s == 3 ? v : d[3];
            }
        } else if (warp_cols) {
            const int i0 = i_lvl - tid;
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int st = s % kScanStages;
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_mbar[warp][st]);
                const unsigned parity = (unsigned)((s / kScanStages) & 1);
                unsigned ok;
                do {
                    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
                } while (!ok);
                const float *col = reinterpret_cast<const float *>(s_dyn + (size_t)(warp * kScanStages + st) * kScanSideBytes) + lane;
                float z[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) z[j] = col[j * 32];
                if (s + kScanStages < 4) {           // the stage is free again: refill it with side s + kScanStages
                    __syncwarp();
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    if (lane == 0) scan_issue_side(s_dyn, &s_mbar[warp][st], &M.lvl[lvl], warp, st, s + kScanStages, i0, b * L.C);
                }
                float m2, inv_s;
                d[s] = dfl_stats16(z, m2, inv_s);
            }
        }
        YS_BSTAMP(2);
        if (live) {
            const int w = L.w[lvl];
            const float st = L.stride[lvl];
            const int iy = i_lvl / w, ix = i_lvl - iy * w;
            const float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
            // dist2bbox in grid units, then * stride (losses/yolo_v8_loss.py:558-559, :412)
            p.x = YS_MUL(YS_SUB(gx, d[0]), st);
            p.y = YS_MUL(YS_SUB(gy, d[1]), st);
            p.z = YS_MUL(YS_ADD(gx, d[2]), st);
            p.w = YS_MUL(YS_ADD(gy, d[3]), st);
            ax = YS_MUL(gx, st);
            ay = YS_MUL(gy, st);
            io.pbox[(size_t)b * L.A + a] = p;
        }
    } else if (live) {
        p = io.pd_box[(size_t)b * L.A + a];
        const float2 c = io.anc[a];
        ax = c.x;
        ay = c.y;
    }
    if (early) {               // k_pack_gt's outputs from here on
        cudaGridDependencySynchronize();
        gt_range<FUSED>(io, b, g0, n);
        ntiles = (n + kGtTile - 1) / kGtTile;
        one_tile = ntiles <= 1;
        tn0 = ntiles > 0 ? min(kGtTile, n) : 0;
        if (tid < tn0) q_pre = io.gt_box[g0 + tid];
    }
    if (tid < tn0) {
        S.box[tid] = q_pre;
        S.area[tid] = area_eps_h(q_pre.x, q_pre.y, q_pre.z, q_pre.w);
        S.any[tid] = 0;
        S.own[tid] = 0;
        S.rkey[tid] = 0ull;
        S.valid[tid] = v_pre ? 1 : 0;
    }
    __syncthreads();
    YS_BSTAMP(3);
    const float area_p = area_eps_h(p.x, p.y, p.z, p.w);
    // bounding box of this warp's anchor centres (non-live lanes neutral); order-preserving float <-> uint map so
    // that the dense entry point may be given any (also negative) anchor points
    // (fused: the centres are positive, and positive floats order like their bit patterns -- no mapping needed)
    float wx_min, wx_max, wy_min, wy_max;
    if (FUSED) {
        wx_min = __uint_as_float(__reduce_min_sync(0xffffffffu, live ? __float_as_uint(ax) : 0x7f800000u));
        wx_max = __uint_as_float(__reduce_max_sync(0xffffffffu, live ? __float_as_uint(ax) : 0u));
        wy_min = __uint_as_float(__reduce_min_sync(0xffffffffu, live ? __float_as_uint(ay) : 0x7f800000u));
        wy_max = __uint_as_float(__reduce_max_sync(0xffffffffu, live ? __float_as_uint(ay) : 0u));
    } else {
        wx_min = ord_to_float(__reduce_min_sync(0xffffffffu, live ? float_to_ord(ax) : 0xffffffffu));
        wx_max = ord_to_float(__reduce_max_sync(0xffffffffu, live ? float_to_ord(ax) : 0u));
        wy_min = ord_to_float(__reduce_min_sync(0xffffffffu, live ? float_to_ord(ay) : 0xffffffffu));
        wy_max = ord_to_float(__reduce_max_sync(0xffffffffu, live ? float_to_ord(ay) : 0u));
    }

    int cnt = 0, first = -1, best_g = 0;
    float best_ov = -1.f;

    // Warp-level prefilter of pass 0: lane j asks whether GT j can contain ANY of the warp's anchor centres (their
    // bounding box); the exact per-anchor test then runs over the few GT that pass, in ascending order.
    auto pass0_tile = [&](int t0, int tn) {
        for (int j0 = 0; j0 < tn; j0 += 32) {
            const int jl = j0 + lane;
            bool rel = false;
            if (jl < tn && S.valid[jl]) {
                const float4 q = S.box[jl];
                rel = q.x < wx_max && q.z > wx_min && q.y < wy_max && q.w > wy_min;
            }
            unsigned m = __ballot_sync(0xffffffffu, rel);
            while (m) {
                const int j = j0 + __ffs(m) - 1;
                m &= m - 1;
                if (live && centre_in_box(ax, ay, S.box[j]) && sel_ok<FUSED>(io, g0 + t0 + j, a)) {
                    if (first < 0) first = t0 + j;
                    ++cnt;
                    S.any[j] = 1;
                }
            }
        }
    };
    if (FUSED) {
        // ONE pass over the GT tiles.  After the centre-in-box test (:210-224) every GT within reach of this warp's
        // anchors -- a decoded box extends at most kDflReach cells from its centre, so IoU(anchor, GT) is exactly 0
        // outside that window -- gets its IoU evaluated once, serving both
        //   * the anchor's own arg-max over all GT slots (:226-244, used if it sits in more than one GT), and
        //   * the GT's arg-max over all anchors (:136, :156, what the forced-assignment loops ask for): redux.max on
        //     the IoU bits (>= 0, so they order as integers), lowest lane among the ties, one shared atomicMax per warp.
        //     Keys with IoU 0 are never recorded: k_pack_gt preset every key to (IoU 0, anchor 0), the first maximum
        //     of an all-zero column.
#ifdef YS_REACH_WINDOW       // r01/r02: the a-priori window (a decoded side is at most kDflReach cells long)
        const float reach = kDflReach * L.stride[lvl];
        const float rx0 = wx_min - reach, rx1 = wx_max + reach, ry0 = wy_min - reach, ry1 = wy_max + reach;
#else
        // "Within reach" of this warp = intersecting the union of its 32 decoded boxes: a GT with q.x >= max p.z (or any of
        // the three symmetric cases) has min(p.z, q.z) - max(p.x, q.x) <= 0 for every lane, and iou_assign_lazy returns
        // exactly 0 for those -- the same floats, no slack needed.  Tighter than the a-priori bound of kDflReach cells
        // around the centres whenever the predicted sides are shorter than 15.5 cells (always, in practice).
        const float rx0 = ord_to_float(__reduce_min_sync(0xffffffffu, live ? float_to_ord(p.x) : 0xffffffffu));
        const float ry0 = ord_to_float(__reduce_min_sync(0xffffffffu, live ? float_to_ord(p.y) : 0xffffffffu));
        // x2 = (gx + d) * stride > 0 and likewise y2: bit patterns order like the values; x1 / y1 may be negative
        const float rx1 = __uint_as_float(__reduce_max_sync(0xffffffffu, live ? __float_as_uint(p.z) : 0u));
        const float ry1 = __uint_as_float(__reduce_max_sync(0xffffffffu, live ? __float_as_uint(p.w) : 0u));
#endif
        const bool warp_live = __any_sync(0xffffffffu, live);
        for (int t = 0; t < ntiles; ++t) {
            const int t0 = t * kGtTile;
            const int tn = min(kGtTile, n - t0);
            if (t > 0) {                                          // tile 0 was staged above
                __syncthreads();
                stage_gt(t0, tn);
                __syncthreads();
            }
            pass0_tile(t0, tn);
            if (warp_live) {
                for (int j0 = 0; j0 < tn; j0 += 32) {
                    const int jl = j0 + lane;
                    bool rel = false;
                    if (jl < tn) {
                        const float4 q = S.box[jl];
                        rel = q.x < rx1 && q.z > rx0 && q.y < ry1 && q.w > ry0;
                    }
                    unsigned m = __ballot_sync(0xffffffffu, rel);
                    while (m) {
                        const int j = j0 + __ffs(m) - 1;
                        m &= m - 1;
                        const float ovr = live ? iou_assign_lazy(p, area_p, S.box[j], S.area[j]) : 0.f;
                        if (ovr > best_ov) { best_ov = ovr; best_g = t0 + j; }    // first maximum (ascending j)
                        const unsigned bits = __float_as_uint(ovr);
                        const unsigned mx = __reduce_max_sync(0xffffffffu, bits);
                        // the block's key for this GT only grows: a warp whose best IoU is below the key's IoU word (a
                        // possibly stale read -- then the atomic below settles it) cannot change it
                        if (mx != 0u && mx >= reinterpret_cast<volatile unsigned *>(&S.rkey[j])[1]) {
                            const unsigned who = __ballot_sync(0xffffffffu, bits == mx);
                            if (lane == __ffs(who) - 1) atomicMax(&S.rkey[j], argmax_key(ovr, (unsigned)a));
                        }
                    }
                }
            }
            if (!one_tile) {
                __syncthreads();
                for (int j = tid; j < tn; j += kScanThreads) {
                    if (S.any[j]) io.gt_flags[g0 + t0 + j] = 1;
                    if (S.rkey[j]) atomicMax(&io.best_key[g0 + t0 + j], S.rkey[j]);
                }
            }
        }
        // all slots outside the window (and possibly some inside) have IoU 0: the first maximum of an all-zero row is slot 0
        if (!(best_ov > 0.f)) { best_ov = 0.f; best_g = 0; }
    } else {
        // dense entry point: arbitrary predicted boxes, no reach bound.  Pass 0: centre-in-box; pass 1: IoU arg-max over
        // all GT slots for anchors inside more than one GT (the forced-assignment arg-max is left to the fix-up).
        for (int pass = 0; pass < 2; ++pass) {
            for (int t = 0; t < ntiles; ++t) {
                const int t0 = t * kGtTile;
                const int tn = min(kGtTile, n - t0);
                if ((pass == 0 && t > 0) || (pass == 1 && !one_tile)) {       // tile 0 of pass 0 was staged above
                    __syncthreads();
                    stage_gt(t0, tn);
                    __syncthreads();
                }
                if (pass == 0) {
                    pass0_tile(t0, tn);
                    if (!one_tile) {
                        __syncthreads();
                        for (int j = tid; j < tn; j += kScanThreads)
                            if (S.any[j]) io.gt_flags[g0 + t0 + j] = 1;
                    }
                } else if (cnt > 1) {
                    for (int j = 0; j < tn; ++j) {
                        const float ov = pair_ov<FUSED>(io, p, area_p, S.box[j], S.area[j]);
                        if (ov > best_ov) { best_ov = ov; best_g = t0 + j; }     // first maximum over ALL slots (:233)
                    }
                }
            }
            // pass 1 only where some anchor sits in more than one GT: per warp when the GT are already staged,
            // per block when pass 1 has to re-stage tiles
            if (pass == 0) {
                if (one_tile) { if (!__any_sync(0xffffffffu, cnt > 1)) break; }
                else if (!__syncthreads_or(cnt > 1)) break;
            }
        }
    }

    YS_BSTAMP(4);
    // ---- preliminary per-anchor resolve (no forced entries yet): losses/yolo_v8_loss.py:226-244
    int g = -1;
    float ov = 0.f;
    if (cnt > 1) { g = best_g; ov = best_ov; }
    else if (cnt == 1) {
        g = first;
        const float4 q = one_tile ? S.box[first] : io.gt_box[g0 + first];
        ov = pair_ov<FUSED, false>(io, p, area_p, q, area_eps_h(q.x, q.y, q.z, q.w));
    }
    float nrm = 0.f;
    int lb = -1;
    if (g >= 0) {
        if (one_tile) atomicAdd(&S.own[g], 1);
        else atomicAdd(&io.owned_cnt[g0 + g], 1);
        lb = label_of<FUSED>(io, g0 + g);
        const float sc = score_at<FUSED>(io, L, b, a, lb);
        nrm = align_norm(sc, ov, io.alpha, io.beta, io.eps);
    }
    if (live) {
        io.tgi[(size_t)b * L.A + a] = g;
        io.norm[(size_t)b * L.A + a] = nrm;
        if (io.lab) io.lab[(size_t)b * L.A + a] = lb;
    }
    // ---- block partial of sum(norm), fixed order; per-GT facts of this block
    float part = warp_sum(nrm);
    if (lane == 0) s_red[warp] = part;
    __syncthreads();
    if (one_tile) {
        for (int j = tid; j < n; j += kScanThreads) {
            if (S.own[j]) atomicAdd(&io.owned_cnt[g0 + j], S.own[j]);
            if (S.any[j]) io.gt_flags[g0 + j] = 1;
            if (FUSED && S.rkey[j]) atomicMax(&io.best_key[g0 + j], S.rkey[j]);
        }
    }
    if (tid == 0) {
        float tsum = 0.f;
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; ++w) tsum += s_red[w];
        io.partial[(size_t)b * gridDim.x + blockIdx.x] = tsum;
    }

    // ---- last block of this image runs the fix-up
    YS_BSTAMP(5);
    // release pattern of a grid barrier: block barrier first, then ONE thread fences (cumulative over the writes the
    // barrier ordered before it) and arrives -- 255 fewer MEMBARs per block than fencing in every thread
    __syncthreads();
    YS_BSTAMP(6);
    if (tid == 0) {
#ifdef YS_ARRIVE_SC
        __threadfence();
#else
        asm volatile("fence.acq_rel.gpu;" ::: "memory");      // release of the block's writes (the barrier ordered them before
                                                              // this thread); lighter than the sequentially consistent membar
#endif
        const unsigned prev = atomicAdd(&io.img_done[b], 1u);
        s_flag = (prev == gridDim.x - 1u);
    }
    __syncthreads();
    YS_BSTAMP(7);
#ifdef YS_SCANTIME
    if (tid == 0 && (blockIdx.x + blockIdx.y * gridDim.x) % 23 == 0)
        printf("BLK x=%d y=%d n=%d t0=%llu sync=%llu tma=%llu dec=%llu scan=%llu res=%llu fence=%llu arr=%llu\n", (int)blockIdx.x,
               (int)blockIdx.y, n, bst[0], bst[1] - bst[0], bst[2] - bst[1], bst[3] - bst[2], bst[4] - bst[3], bst[5] - bst[4],
               bst[6] - bst[5], bst[7] - bst[6]);
#endif
    if (!s_flag) return;
    __threadfence();
    if (io.skip_fixup) {
        if (tid == 0) io.img_done[b] = 0u;
        return;
    }
    // fix-up scratch: fused -- the 64 KB tile (consumed by the decode above); dense -- the scan's own shared memory
    if (FUSED) fixup_image<FUSED>(L, io, b, B, s_dyn, kFixSmemGtFused, (int)gridDim.x);
    else fixup_image<FUSED>(L, io, b, B, &sm.f, kFixSmemGt, (int)gridDim.x);
}

// ---------------------------------------------------------------------------------------------------
// Fix-up of one image (all kScanThreads threads of the last block to finish that image).
// Phases, separated by block barriers; everything that is parallel in the reference's semantics runs on all
// eight warps, only the order-dependent bookkeeping of loop 2 is serial (one thread, shared-memory state):
//   A  classify the GT: lonely (no anchor centre inside -> loop 1 fires) / ownerless (loop 2 may fire)
//   B  arg-max over all anchors of IoU(anchor, GT) for those GT (block-wide, first maximum)
//   1  first resolve (:142) of the anchors loop 1 forced, one warp per anchor
//   B' (rare) arg-max for GT that lost their only anchor in phase 1
//   2  loop 2 (:146-162), sequential in g.  Its tgi/fg edits are discarded by the final resolve (:165), so only
//      the per-GT owner counts are tracked: gold = current owner of the stolen anchor.
//   3  final resolve from mask_pos alone for every anchor carrying a forced entry, one warp per anchor
// ---------------------------------------------------------------------------------------------------
template <bool FUSED>
__device__ __forceinline__ void fix_argmax(const LevelSet &L, const AssignIO &io, int b, int g0, const int *list,
                                           int nn, unsigned long long *keyp) {
    const int tid = threadIdx.x, lane = tid & 31;
    const float4 *pb = (FUSED ? io.pbox : io.pd_box) + (size_t)b * L.A;
    for (int q = 0; q < nn; ++q) {
        const int j = list[q];
        const float4 gq = io.gt_box[g0 + j];
        const float area_q = area_eps_h(gq.x, gq.y, gq.z, gq.w);
        float best = -1.f;          // thread-local first maximum over this thread's anchors (ascending a)
        int best_a = 0;
        // fused: the per-anchor state lives in the padded anchor space and only the live entries were written.
        // Four boxes are fetched before the first is used (the loop was one L2 round trip per anchor otherwise: 12 us
        // for the 8400 anchors of a 640 px image, measured with -DYS_FIXTIME).
        for (int l = 0; l < (FUSED ? L.n : 1); ++l) {
            const int a0 = FUSED ? L.a_off[l] : 0, a1 = a0 + (FUSED ? L.hw[l] : L.A);
            for (int ab = a0 + tid; ab < a1; ab += 4 * kScanThreads) {
                float4 pv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int a = ab + u * kScanThreads;
                    pv[u] = a < a1 ? (FUSED ? __ldcg(pb + a) : pb[a]) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int a = ab + u * kScanThreads;
                    if (a < a1) {
                        const float ov = pair_ov<FUSED>(io, pv[u], area_eps_h(pv[u].x, pv[u].y, pv[u].z, pv[u].w), gq, area_q);
                        if (ov > best) { best = ov; best_a = a; }
                    }
                }
            }
        }
        unsigned long long key = best >= 0.f ? argmax_key(best, (unsigned)best_a) : 0ull;
        key = warp_max(key);
        if (lane == 0) atomicMax(&keyp[j], key);
    }
}

#ifdef YS_FIXTIME         // timing experiment: thread 0 of every fix-up prints its phase stamps (ns) once it is done
#define YS_STAMP(i) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); stamp[i] = t_; } } while (0)
#else
#define YS_STAMP(i) do { } while (0)
#endif
#ifndef YS_FIXDIAG
#define YS_FIXDIAG 0      // timing experiments only: bit0 skip phase 1, bit1 skip loop 2, bit2 skip phase 3 (results wrong)
#endif
template <bool FUSED>
__device__ void fixup_image(const LevelSet &L, const AssignIO &io, int b, int B, void *scratch, int cap, int nblk) {
    __shared__ int s_nneedy, s_ncmp, s_nlost, s_resume, s_nfired, s_want, s_nuniq, s_nlonely;
    __shared__ double s_delta[kScanThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kScanThreads / 32;
    int g0, n;
    gt_range<FUSED>(io, b, g0, n);
    volatile int *tgi = io.tgi + (size_t)b * L.A;      // written by other blocks of this launch: no L1
    int *labp = io.lab ? io.lab + (size_t)b * L.A : nullptr;
    volatile float *norm = io.norm + (size_t)b * L.A;
    // per-GT scratch: shared memory when the image's GT fit, else the global scratch (8 ints per GT) in place
    const bool in_smem = n <= cap;
    unsigned long long *s_key = reinterpret_cast<unsigned long long *>(scratch);      // [cap], then 7 int arrays of cap
    int *s_int = reinterpret_cast<int *>(s_key + cap);
    int *gs = io.forced + 8 * (size_t)g0;
    int *cntp = in_smem ? s_int : (io.owned_cnt + g0);
    int *forced = in_smem ? s_int + cap : gs;                               // [2j] loop-1 anchor, [2j+1] loop-2 anchor
    int *needy = in_smem ? s_int + 3 * cap : (gs + 2 * (size_t)n);          // GT whose arg-max anchor is needed
    int *kk = in_smem ? s_int + 4 * cap : (gs + 3 * (size_t)n);             // arg-max anchor of GT j (-1: not computed)
    int *tg0 = in_smem ? s_int + 5 * cap : (gs + 4 * (size_t)n);            // owner GT of anchor kk[j] after phase 1
    int *lost = in_smem ? s_int + 6 * cap : (gs + 5 * (size_t)n);
    unsigned long long *keyp = in_smem ? s_key : (io.best_key + g0);

#ifdef YS_FIXTIME
    unsigned long long stamp[8] = {};
#endif
    YS_STAMP(0);
    // the image's scan partials (independent of the fix-up): warp 0 fetches them now, their latency hides under phase A
    double acc = 0.0;
    if (warp == 0)
        for (int i = lane; i < nblk; i += 32) acc += (double)__ldcg(io.partial + (size_t)b * nblk + i);
    if (tid == 0) { s_nneedy = 0; s_ncmp = 0; s_nlost = 0; s_nuniq = 0; s_nlonely = 0; s_nfired = 0; }
    if (tid < kWarps) s_delta[tid] = 0.0;
    __syncthreads();
    // ---- A: which GT need a forced assignment?  (common case: none)
    for (int j = tid; j < n; j += kScanThreads) {
        const bool valid = FUSED ? true : (io.gt_valid[g0 + j] != 0.f);
        const int c = __ldcg(io.owned_cnt + g0 + j);
        const int fl = __ldcg(io.gt_flags + g0 + j);
        if (in_smem) cntp[j] = c;
        const bool lonely1 = valid && fl == 0;           // loop 1 (:117-139): valid GT with no anchor centre inside
        forced[2 * j] = lonely1 ? -2 : -1;               // -2: pending, resolved below
        if (lonely1) atomicAdd(&s_nlonely, 1);
        forced[2 * j + 1] = -1;
        // risk GT: the scan already found the arg-max anchor (all blocks of the image are done: key is final)
        const bool have = FUSED;      // fused: the scan left every GT's arg-max key in best_key (all blocks are done)
        const unsigned long long key = have ? __ldcg(io.best_key + g0 + j) : 0ull;
        if (in_smem || !have) keyp[j] = key;
        kk[j] = have ? (int)argmax_key_index(key) : -1;
        if (valid && (lonely1 || c == 0)) {              // loop 2 (:146-162) can only fire for GT that own nothing now
            needy[atomicAdd(&s_nneedy, 1)] = j;
            if (!have) lost[atomicAdd(&s_ncmp, 1)] = j;  // lost[] doubles as the "arg-max still to compute" list here
        }
    }
    __syncthreads();
    const int nneedy = s_nneedy;
    YS_STAMP(1);
    if (nneedy > 0) {      // block-uniform
        // ---- B: arg-max over all anchors of IoU(anchor, GT j), first maximum (:136, :156), for the needy GT
        //      the scan did not cover
        if (s_ncmp > 0) fix_argmax<FUSED>(L, io, b, g0, lost, s_ncmp, keyp);
        __syncthreads();
        for (int q = tid; q < nneedy; q += kScanThreads) {
            const int j = needy[q];
            const int k = (int)argmax_key_index(keyp[j]);
            kk[j] = k;
            if (forced[2 * j] == -2) forced[2 * j] = k;
        }
        __syncthreads();
        // ---- 1: first resolve (:142) for the anchors loop 1 touched; everyone else keeps the scan's result
        for (int q = warp; q < nneedy; q += kWarps) {
            if (YS_FIXDIAG & 1) break;
            const int j = needy[q];
            const int k = forced[2 * j];
            if (k < 0) continue;
            bool seen = false;   // k already handled through another lonely GT?  (lowest j owns the job)
            for (int jj = lane; jj < n; jj += 32) seen |= (jj < j && forced[2 * jj] == k);
            if (__any_sync(0xffffffffu, seen)) continue;
            float ov;
            const int gnew = resolve_row<FUSED>(io, L, b, k, g0, n, forced, false, ov);
            if (lane == 0) {
                const int gold = tgi[k];
                if (gold >= 0) atomicSub(&cntp[gold], 1);
                if (gnew >= 0) atomicAdd(&cntp[gnew], 1);
                tgi[k] = gnew;
            }
        }
        __syncthreads();
        YS_STAMP(2);
        // ---- B': GT that own nothing now but were not on the needy list (they lost their only anchor in phase 1)
        for (int j = tid; j < n; j += kScanThreads) {
            const bool valid = FUSED ? true : (io.gt_valid[g0 + j] != 0.f);
            if (valid && cntp[j] == 0 && kk[j] < 0) lost[atomicAdd(&s_nlost, 1)] = j;
        }
        __syncthreads();
        const int nlost = s_nlost;
        if (nlost > 0) {
            fix_argmax<FUSED>(L, io, b, g0, lost, nlost, keyp);
            __syncthreads();
            for (int q = tid; q < nlost; q += kScanThreads) kk[lost[q]] = (int)argmax_key_index(keyp[lost[q]]);
            __syncthreads();
        }
        // owner of every candidate anchor after phase 1
        for (int j = tid; j < n; j += kScanThreads) tg0[j] = kk[j] >= 0 ? tgi[kk[j]] : -1;
        __syncthreads();
        // ---- 2: loop 2 (:146-162): sequential in g; later GT see the edits of earlier ones.  Warp 0 walks the GT
        //      32 at a time (ballot over "owns nothing"), handles the first such GT and looks again from the next
        //      one on, because the anchor it just took may have been the only one of a later GT.  The GT that
        //      fired and their anchors are listed in lost[] / needy[] (needy[] is free after phase 1): "who owns
        //      anchor k now" = the latest listed thief of k, else the owner after phase 1.
        //      Dense path only: a GT whose only anchor is stolen inside this loop has no arg-max yet -- the walk
        //      stops there, the block computes it, and the walk resumes at the same GT.
        YS_STAMP(3);
        int *thief_k = needy;
        if (tid == 0) { s_resume = 0; s_nfired = 0; s_want = -1; }
        __syncthreads();
        while (true) {
            if (warp == 0 && !(YS_FIXDIAG & 2)) {
                int nf = s_nfired, want = -1, c = s_resume;
                while (c < n) {
                    const int jj = c + lane;
                    const bool z = jj < n && cntp[jj] == 0 && (FUSED || io.gt_valid[g0 + jj] != 0.f);
                    const unsigned m = __ballot_sync(0xffffffffu, z);
                    if (m == 0u) { c += 32; continue; }
                    const int jf = c + __ffs(m) - 1;
                    const int k = kk[jf];
                    if (k < 0) { want = jf; break; }
                    int latest = -1;
                    for (int f0 = 0; f0 < nf; f0 += 32) {
                        const int f = f0 + lane;
                        const unsigned hm = __ballot_sync(0xffffffffu, f < nf && thief_k[f] == k);
                        if (hm) latest = f0 + 31 - __clz(hm);
                    }
                    if (lane == 0) {
                        const int gold = latest >= 0 ? lost[latest] : tg0[jf];
                        if (gold >= 0) cntp[gold] -= 1;
                        cntp[jf] += 1;
                        forced[2 * jf + 1] = k;
                        lost[nf] = jf;
                        thief_k[nf] = k;
                    }
                    ++nf;
                    __syncwarp();
                    c = jf + 1;
                }
                if (lane == 0) { s_nfired = nf; s_want = want; s_resume = want >= 0 ? want : n; }
            }
            __syncthreads();
            if (FUSED || s_want < 0) break;                     // block-uniform
            fix_argmax<FUSED>(L, io, b, g0, &s_want, 1, keyp);
            __syncthreads();
            if (tid == 0) {
                const int k = (int)argmax_key_index(keyp[s_want]);
                kk[s_want] = k;
                tg0[s_want] = tgi[k];                           // loop 2 itself never writes tgi
            }
            __syncthreads();
        }
        // ---- 3: final resolve (:165) from mask_pos alone, for every anchor that carries a forced entry.
        //      3a  the anchors to redo, listed in a fixed order: loop 1's in GT order (ballot compaction by warp 0),
        //          then loop 2's in firing order (they are thief_k[] already).  A fixed order keeps the fp64 sum
        //          of the target-score corrections the same every run.
        //      3b  one owner per distinct anchor: the first row of the list naming it
        //      3c  rows resolved by groups of W lanes, W from the row count; the group's first lane finishes
        //          the row (label, score, norm), so those global round trips overlap across the block
        YS_STAMP(4);
        const int nf = (YS_FIXDIAG & 4) ? 0 : s_nfired;
        int *row_all = tg0;          // [n1 + nf] <= 2n ints: tg0[] + lost[] (the thief list is not needed as GT any more)
        int *row_k = needy;          // [n1 + nf] ints: needy[] + kk[]; -1 = a duplicate of an earlier row
        if (warp == 0) {
            int base = 0;
            if (s_nlonely > 0 && !(YS_FIXDIAG & 4)) {
                for (int c = 0; c < n; c += 32) {
                    const int jj = c + lane;
                    const int k = jj < n ? forced[2 * jj] : -1;
                    const unsigned m = __ballot_sync(0xffffffffu, k >= 0);
                    if (k >= 0) row_all[base + __popc(m & ((1u << lane) - 1u))] = k;
                    base += __popc(m);
                }
            }
            if (lane == 0) s_nuniq = base;
        }
        __syncthreads();
        const int n1 = s_nuniq;
        const int nrow = n1 + nf;
        for (int f = tid; f < nf; f += kScanThreads) row_all[n1 + f] = thief_k[f];      // needy[] -> tg0[]/lost[]: disjoint
        __syncthreads();
        for (int q = tid; q < nrow; q += kScanThreads) {
            const int k = row_all[q];
            bool dup = false;
#pragma unroll 4
            for (int q2 = 0; q2 < q; ++q2) dup |= (row_all[q2] == k);
            row_k[q] = dup ? -1 : k;
        }
        __syncthreads();
        YS_STAMP(5);
        int W = 32;
        while (W > 1 && nrow * W > kScanThreads) W >>= 1;
        const int groups = kScanThreads / W;
        double delta = 0.0;
        for (int q0 = 0; q0 < nrow; q0 += groups) {
            const int q = q0 + tid / W;
            const int k = q < nrow ? row_k[q] : -1;
            const bool active = k >= 0;
            float ov;
            const int gfin = resolve_row_group<FUSED>(io, L, b, k, g0, n, forced, W, active, ov);
            if (active && (tid & (W - 1)) == 0) {
                tgi[k] = gfin;
                float nrm = 0.f;
                int lb = -1;
                if (gfin >= 0) {
                    lb = label_of<FUSED>(io, g0 + gfin);
                    const float sc = score_at<FUSED>(io, L, b, k, lb);
                    nrm = align_norm(sc, ov, io.alpha, io.beta, io.eps);
                }
                delta += (double)nrm - (double)norm[k];
                norm[k] = nrm;
                if (labp) labp[k] = lb;
            }
        }
        delta = warp_sum(delta);
        YS_STAMP(6);
        if (lane == 0) s_delta[warp] = delta;
    }
    __syncthreads();
#ifdef YS_FIXTIME
    if (tid == 0)
        printf("FIX b=%d y=%d n=%d needy=%d uniq=%d t0=%llu A=%llu p1=%llu Bp=%llu l2=%llu fl=%llu p3=%llu\n", b, (int)blockIdx.y, n,
               nneedy, s_nuniq + s_nfired, stamp[0], stamp[1] - stamp[0], stamp[2] - stamp[1], stamp[3] - stamp[2], stamp[4] - stamp[3],
               stamp[5] - stamp[4], stamp[6] - stamp[5]);
#endif
    // ---- leave the per-GT scratch zeroed for the next call
    for (int j = tid; j < n; j += kScanThreads) {
        io.owned_cnt[g0 + j] = 0;
        io.gt_flags[g0 + j] = 0;
    }
    // ---- per-image sum of target scores (fixed order), then target_scores_sum in the last image
    if (warp == 0) {
        acc = warp_sum(acc);
        unsigned prev = 0u;
        if (lane == 0) {
            double dsum = 0.0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) dsum += s_delta[w];
            io.img_sum[b] = acc + dsum;
            io.img_done[b] = 0u;
            __threadfence();
            prev = atomicAdd(io.done_counter, 1u);
        }
        prev = __shfl_sync(0xffffffffu, prev, 0);
        if (prev == (unsigned)B - 1u) {                // last image of the batch: target_scores_sum, lanes over the images
            __threadfence();
            double tot = 0.0;
            for (int i = lane; i < B; i += 32) tot += __ldcg(io.img_sum + i);
            tot = warp_sum(tot);
            if (lane == 0) {
                const float tss = (float)tot;
                io.tss[0] = tss > 1.f ? tss : 1.f;     // max(target_scores.sum(), 1)  (:472)
                *io.done_counter = 0u;
            }
        }
    }
}

}  // namespace ys
