// ys_loss_stream.cuh -- k_loss_stream: the loss forward + backward kernel as a persistent, TMA-fed stream.
//
// Same arithmetic as k_loss (ys_loss.cu) -- BCE-with-logits, CIoU and DFL, forward and backward, replacing
// losses/yolo_v8_loss.py:472-499 and :280-324 of the reference plus their autograd backward -- but the head values
// are no longer loaded into registers by the warp that consumes them.  k_loss is latency bound (64 registers x
// 32 warps per SM leave ~40 KB of loads in flight per SM); here every warp owns a two- or three-stage ring of 4-5 KB
// shared-memory tiles filled by bulk asynchronous copies (cp.async.bulk global -> shared, completion on an
// mbarrier), so the tile of work item k+1 is in flight -- costing no registers -- while item k is computed:
// 16 warps x 5 KB = 80 KB in flight per SM at all times.
//
//   grid      one CTA per SM (persistent), 12 warps (measured best of 8..20); rounds over contiguous windows of groups of 12 items
//   items     image-major, per image first the box items, then the class items:
//     box     16 anchors x 64 box channels (64 rows of 64 B).  Lane = (side pair, anchor): each half-warp runs two
//             of the four sides, both halves evaluate the anchor's CIoU.  Items without a foreground anchor load
//             nothing and only zero-fill their 4 KB of gradient.
//     class   128 anchors x kClsRows (10) class channels (rows of 512 B), float4 per lane per channel; 10 rows measured
//             1.5 % faster than 8 (fewer items, same registers), 16 no better.
//   aux       the per-anchor assignment the item needs (labels + target scores, or decoded boxes + target scores)
//             rides in the same stage through the same mbarrier.
//   output    gradients go straight from registers to HBM (full 32-byte sectors per store instruction).
// Per-warp partial sums are carried across items and reduced once per CTA; the last CTA reduces the grid's
// partials in a fixed order (fp64).  The item -> warp map is static, so results are run-to-run deterministic.
#pragma once
#include <cuda.h>
#include "ys_common.cuh"

namespace ys {

#ifndef YS_STREAM_WARPS
#define YS_STREAM_WARPS 12
#endif
constexpr int kStreamWarps = YS_STREAM_WARPS;
constexpr int kStreamThreads = kStreamWarps * 32;
constexpr int kStreamStagesMax = 3;      // tiles per warp ring (template parameter STAGES = 2 or 3): one being computed, the others in flight
constexpr int kBoxItem = 16;             // anchors per box item
#ifndef YS_CLS_ROWS
#define YS_CLS_ROWS 10
#endif
constexpr int kClsRows = YS_CLS_ROWS;    // class channels per class item
constexpr int kTileFloats = (kClsRows * 128 > 1024 ? kClsRows * 128 : 1024) + 16;   // 4 KB tile; the second side pair of a box tile sits 16 floats further (bank spread)
constexpr int kAuxBytes = 1024;
constexpr int kStageBytes = kTileFloats * 4 + kAuxBytes + 64;   // 5248, a multiple of 128
static_assert(kStageBytes % 128 == 0, "stage size");

struct StreamParams {
    const int *tgi;          // (B,A) -1 = background
    const int *lab;          // (B,A) class of the assigned GT, -1 = background
    const float *norm;       // (B,A)
    const int *gt_off;       // (B+1)
    const float4 *gt_box;    // px
    const float4 *pbox;      // (B,A) decoded boxes, px (written by k_assign)
    const float *tss;        // [0]
    float4 *partials;        // (gridDim.x)
    unsigned *done;          // arrival counter, self-resetting
    float *loss_out;         // [4]
    int nc, n_groups;        // class channels, ceil(nc / kClsRows)
    int n_box;               // box items per image
    int items_per_img;       // n_box + n_chunks * n_groups
    int total_items;         // B * items_per_img
    unsigned n_groups16;     // groups of kStreamWarps consecutive items (one item per warp of a CTA)
    unsigned rounds;         // ceil(n_groups16 / gridDim.x): every CTA runs this many rounds (groups past the end are padding)
    unsigned ipi_magic, nch_magic;   // floor(2^32 / items_per_img), floor(2^32 / n_chunks): division by multiply-high
    unsigned rotate;         // per-round rotation of the CTA -> group map inside the round's window (see the kernel)
    float hyp_box, hyp_cls, hyp_dfl;
    float batch;
};

// One 2-D tensor map per level and item kind over the level viewed as [B*C rows][h*w columns] fp32: a whole tile is
// ONE TMA operation (row-by-row 1-D bulk copies of 64..512 B are TMA issue-rate bound: measured 2.8 TB/s).
struct StreamMaps {
    CUtensorMap box[YS_MAX_LEVELS];     // box {16 columns, 64 rows}
    CUtensorMap cls[YS_MAX_LEVELS];     // box {128 columns, 8 rows}
};

// Gradient stores of the streamed (written once, never re-read here) rows.  YS_STREAM_ST: 0 plain, 1 st.global.cs
// (evict-first), 2 st.global.wt (write-through)
#ifndef YS_STREAM_ST
#define YS_STREAM_ST 0
#endif
#if YS_STREAM_ST == 1
#define YS_ST4(p, v) __stcs((p), (v))
#elif YS_STREAM_ST == 2
#define YS_ST4(p, v) __stwt((p), (v))
#else
#define YS_ST4(p, v) (*(p) = (v))
#endif

// ---- mbarrier / bulk-copy primitives (PTX ISA: mbarrier, cp.async.bulk) ------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tma_g2s_2d(unsigned dst, const CUtensorMap *map, int col, int row, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(col), "r"(row), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- work item ------------------------------------------------------------------------------------------------
struct Item {
    int b, l, hw, i0;        // image, level, level size, first anchor within the level
    int group;               // class items: channel group; box items: -1
    int n_live;              // anchors of the item that exist (<= 16 or <= 128); 0 = padding unit past the last item
};
__device__ __forceinline__ Item decode_item(const LevelSet &L, const StreamParams &P, int item) {
    Item it;
    if (item >= P.total_items) {            // padding unit of the last (partial) group
        it.b = 0; it.l = 0; it.hw = 0; it.i0 = 0; it.group = -1; it.n_live = 0;
        return it;
    }
    // item / items_per_img by multiply-high with floor(2^32 / d) (never too large, at most one too small: one correction)
    int bq = (int)__umulhi((unsigned)item, P.ipi_magic);
    int r = item - bq * P.items_per_img;
    if (r >= P.items_per_img) { ++bq; r -= P.items_per_img; }
    it.b = bq;
    // Items of an image: first the box items, then the class items (see the schedule note in the kernel; interleaving
    // the two kinds in proportion measured 4 % slower).
    int l = 0;
    if (r < P.n_box) {
#pragma unroll
        for (int k = 1; k < YS_MAX_LEVELS; ++k)
            if (k < L.n && r >= L.hseg_off[k]) l = k;
        it.i0 = (r - L.hseg_off[l]) * kBoxItem;
        it.group = -1;
        it.hw = L.hw[l];
        it.n_live = min(kBoxItem, it.hw - it.i0);
    } else {
        const int q = r - P.n_box;
#ifdef YS_CLS_CHUNK_MAJOR
        const int chunk = q / P.n_groups;
        it.group = q - chunk * P.n_groups;
#else
        // group-major: consecutive items (= the warps of one CTA) walk consecutive 512-byte pieces of the same rows
        int gq = (int)__umulhi((unsigned)q, P.nch_magic);
        int chunk = q - gq * L.n_chunks;
        if (chunk >= L.n_chunks) { ++gq; chunk -= L.n_chunks; }
        it.group = gq;
#endif
#pragma unroll
        for (int k = 1; k < YS_MAX_LEVELS; ++k)
            if (k < L.n && chunk >= L.chunk_off[k]) l = k;
        it.i0 = (chunk - L.chunk_off[l]) * kChunk;
        it.hw = L.hw[l];
        it.n_live = min(kChunk, it.hw - it.i0);
    }
    it.l = l;
    return it;
}

// Queue entry of the pipeline: an Item in three registers (meta = (group + 1) << 12 | level << 8 | n_live).
struct ItemQ {
    int b, i0, meta;
};
__device__ __forceinline__ ItemQ pack_item(const Item &it) {
    ItemQ q;
    q.b = it.b; q.i0 = it.i0; q.meta = ((it.group + 1) << 12) | (it.l << 8) | it.n_live;
    return q;
}
__device__ __forceinline__ Item unpack_item(const LevelSet &L, const ItemQ &q) {
    Item it;
    it.b = q.b; it.i0 = q.i0;
    it.n_live = q.meta & 255; it.l = (q.meta >> 8) & 15; it.group = (q.meta >> 12) - 1;
    it.hw = it.n_live > 0 ? L.hw[it.l] : 0;
    return it;
}

// Issue the asynchronous copies of one item into a stage (lane 0 issues; the stage's mbarrier completes when all bytes
// have landed).  Box items: only when the item has a foreground anchor.  The per-anchor state always travels by bulk copy
// (the padded anchor space keeps every item's first entry 16-byte aligned; counts are rounded up to 4 entries, which
// stays inside the level's padding).  The head-value tile comes by ONE tensor operation when the level has a tensor map
// (L.tma); otherwise the consuming warp fills it with plain loads right after the wait (fill_tile_plain).
__device__ __forceinline__ void issue_item(const LevelSet &L, const StreamParams &P, const StreamMaps &M, const Item &it,
                                           char *stage, unsigned bar, int lane) {
    const unsigned tile = smem_u32(stage), aux = tile + kTileFloats * 4;
    const size_t arow = (size_t)it.b * L.A + L.a_off[it.l] + it.i0;
    const unsigned n4 = (unsigned)((it.n_live + 3) & ~3);
    const unsigned rb = n4 * 4u;
    const bool tma = L.tma[it.l] != 0;
    // the tensor operation always delivers the full box (columns past the level's end arrive as zeros)
    if (it.group < 0) {
        if (lane == 0) {
            mbar_expect_tx(bar, (tma ? 4096u : 0u) + (unsigned)it.n_live * 16u + rb);         // tile + pbox (16 B) + norm (4 B)
            if (tma) tma_g2s_2d(tile, &M.box[it.l], it.i0, it.b * L.C, bar);
            bulk_g2s(aux, P.pbox + arow, (unsigned)it.n_live * 16u, bar);
            bulk_g2s(aux + 256u, P.norm + arow, rb, bar);
        }
    } else {
        if (lane == 0) {
            mbar_expect_tx(bar, (tma ? (unsigned)(kClsRows * 512) : 0u) + 2u * rb);
            if (tma) tma_g2s_2d(tile, &M.cls[it.l], it.i0, it.b * L.C + 64 + it.group * kClsRows, bar);
            bulk_g2s(aux, P.lab + arow, rb, bar);
            bulk_g2s(aux + 512u, P.norm + arow, rb, bar);
        }
    }
}

// Plain-load tile loader for levels without a tensor map (h*w % 4 != 0, or base pointers that are not 16-byte aligned):
// the warp copies the item's rows into the stage in the layout the tensor operation would have produced -- box items
// [64 rows][16 anchors], class items [rows][128 anchors] -- every row piece one contiguous, coalesced run; columns past
// the level's end are written as zeros.  Ends on a __syncwarp().
__device__ __forceinline__ void fill_tile_plain(const LevelSet &L, const StreamParams &P, const Item &it, char *stage, int lane) {
    float *tile = reinterpret_cast<float *>(stage);
    const size_t hw = (size_t)it.hw;
    if (it.group < 0) {
        const float *src = L.x[it.l] + (size_t)it.b * L.C * hw + it.i0;
        const int a16 = lane & 15, rsub = lane >> 4;
        const bool live = a16 < it.n_live;
#pragma unroll 8
        for (int r = rsub; r < 64; r += 2) tile[r * kBoxItem + a16] = live ? __ldg(src + (size_t)r * hw + a16) : 0.f;
    } else {
        const int c_lo = it.group * kClsRows;
        const int c_n = min(kClsRows, P.nc - c_lo);
        const float *src = L.x[it.l] + ((size_t)it.b * L.C + 64 + c_lo) * hw + it.i0;
        for (int r = 0; r < c_n; ++r) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = lane + 32 * j;
                tile[r * kChunk + col] = col < it.n_live ? __ldg(src + (size_t)r * hw + col) : 0.f;
            }
        }
    }
    __syncwarp();
}

// ---- box item -------------------------------------------------------------------------------------------------
template <bool WRITE_GRAD>
__device__ __forceinline__ void stream_box_zero(const LevelSet &L, const Item &it, int lane) {
    if (!WRITE_GRAD) return;
    float *g = L.g[it.l] + (size_t)it.b * L.C * it.hw + it.i0;
    if (L.tma[it.l]) {
        // 64 rows x 64 B: 4 lanes per row, 8 rows per instruction (h*w % 4 == 0: the 4 anchors of a piece are live together)
        const int col = (lane & 3) * 4, rsub = lane >> 2;
        if (col < it.n_live) {
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int c = 0; c < 8; ++c) YS_ST4(reinterpret_cast<float4 *>(g + (size_t)(c * 8 + rsub) * it.hw + col), z4);
        }
    } else {
        // rows without 16-byte alignment: 16 lanes per row, 2 rows per instruction, scalar stores
        const int col = lane & 15, rsub = lane >> 4;
        if (col < it.n_live) {
#pragma unroll 8
            for (int c = 0; c < 32; ++c) g[(size_t)(c * 2 + rsub) * it.hw + col] = 0.f;
        }
    }
}

template <bool WRITE_GRAD>
__device__ __forceinline__ void stream_box(const LevelSet &L, const StreamParams &P, const Item &it, const char *stage,
                                           int tg, int lane, float inv_tss, float &acc_box, float &acc_dfl) {
    const float *tile = reinterpret_cast<const float *>(stage);
    const char *aux = stage + kTileFloats * 4;
    const int sp = lane >> 4, a16 = lane & 15;
    const int i = it.i0 + a16;
    const bool live = a16 < it.n_live;
    const bool fg = live && tg >= 0;
    const float st = L.stride[it.l];
    float gd0 = 0.f, gd1 = 0.f, tt0 = 0.f, tt1 = 0.f;     // this lane's two sides: CIoU gradient wrt the distance, DFL target
    float cdfl = 0.f, wgt = 0.f;
    if (fg) {
        wgt = reinterpret_cast<const float *>(aux + 256)[a16];                        // target_scores.sum(-1) (:298)
        const float4 q = __ldg(P.gt_box + __ldg(P.gt_off + it.b) + tg);
        const float4 tb = make_float4(q.x / st, q.y / st, q.z / st, q.w / st);        // target_bboxes /= stride (:479)
        const float4 pp = reinterpret_cast<const float4 *>(aux)[a16];                 // decoded by k_assign, pixels
        const float4 pb = make_float4(pp.x / st, pp.y / st, pp.z / st, pp.w / st);    // back to grid units (:412)
        const int w = L.w[it.l];
        const int iy = i / w, ix = i - iy * w;
        const float gx = (float)ix + 0.5f, gy = (float)iy + 0.5f;
        float g4[4];
        const float ciou = ciou_with_grad(pb, tb, g4);
        if (sp == 0) acc_box += (1.0f - ciou) * wgt;                                  // one half-warp accounts the value
        const float kb = -(P.hyp_box * P.batch * inv_tss) * wgt;                      // L = (1-ciou)*w/tss
        // d/d(dist): x1 = gx - d0 -> -g4[0]; y1 -> -g4[1]; x2 = gx + d2 -> g4[2]; y2 -> g4[3]
        // bbox2dist targets (:304, :327-330): l = gx - x1, t = gy - y1, r = x2 - gx, b = y2 - gy
        if (sp == 0) { gd0 = kb * -g4[0]; gd1 = kb * -g4[1]; tt0 = gx - tb.x; tt1 = gy - tb.y; }
        else         { gd0 = kb * g4[2];  gd1 = kb * g4[3];  tt0 = tb.z - gx; tt1 = tb.w - gy; }
        cdfl = P.hyp_dfl * P.batch * inv_tss * 0.25f * wgt;
    }
    float *gs = WRITE_GRAD ? (L.g[it.l] + (size_t)it.b * L.C * it.hw + i) : nullptr;
    float ce = 0.f;
#pragma unroll 1
    for (int s2 = 0; s2 < 2; ++s2) {
        const int row0 = (sp * 2 + s2) * 16;                                          // first channel row of this side
        const float *zs = tile + row0 * kBoxItem + a16;
        float z[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) z[j] = fg ? zs[j * kBoxItem] : 0.f;
        float m = z[0];
#pragma unroll
        for (int j = 1; j < 16; ++j) m = fmaxf(m, z[j]);
        const float mm = m * kLog2e;
        float ssum = 0.f, dsum = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            z[j] = ex2_fast(fmaf(z[j], kLog2e, -mm));
            ssum += z[j];
            dsum = fmaf(z[j], (float)j, dsum);
        }
        const float inv = rcp_fast(ssum);
        const float d = dsum * inv;
        const float t = fminf(fmaxf(s2 ? tt1 : tt0, 0.f), 14.99f);                    // reg_max - 1 - 0.01 (:304)
        const int tli = (int)t;
        const float wls = (float)(tli + 1) - t, wrs = 1.0f - wls;
        const float zl = fg ? zs[tli * kBoxItem] : 0.f, zr = fg ? zs[(tli + 1) * kBoxItem] : 0.f;
        // cross-entropy at bins tl and tl+1 (:315-324): logsumexp = (mm + log2(sum)) * ln2
        const float lse = (mm + __log2f(ssum)) * 0.69314718055994531f;
        ce += (lse - zl) * wls + (lse - zr) * wrs;
        if (WRITE_GRAD && live) {
            // g_j = p_j * (gd*(j - d) + cdfl) - cdfl * ([j == tl] * wl + [j == tl+1] * wr),  p_j = e_j * inv
            const float gds = s2 ? gd1 : gd0;
            const float ka = inv * gds, kb0 = inv * fmaf(-gds, d, cdfl);
            float *gr = gs + (size_t)row0 * it.hw;
#pragma unroll
            for (int j = 0; j < 16; ++j) gr[(size_t)j * it.hw] = z[j] * fmaf(ka, (float)j, kb0);
            if (fg) {      // the two target bins get their one-hot part (same thread, same address: ordered)
                const float el = ex2_fast(fmaf(zl, kLog2e, -mm)), er = ex2_fast(fmaf(zr, kLog2e, -mm));
                gr[(size_t)tli * it.hw] = fmaf(el, fmaf(ka, (float)tli, kb0), -cdfl * wls);
                gr[(size_t)(tli + 1) * it.hw] = fmaf(er, fmaf(ka, (float)(tli + 1), kb0), -cdfl * wrs);
            }
        }
    }
    acc_dfl += ce * 0.25f * wgt;                                                      // mean over 4 sides (:324) * weight
}

// ---- class item -----------------------------------------------------------------------------------------------
// ALIGNED (levels with 16-byte aligned rows): lane = anchors 4*lane .. 4*lane+3, one float4 per row from shared memory and
// one float4 store per row.  Otherwise lane = anchors lane, lane+32, lane+64, lane+96: scalar shared-memory reads and
// scalar, fully coalesced stores, each anchor live on its own.  Anchors past the level's end read a logit of -inf, which
// contributes exactly zero to every sum (exp(-inf) = 0, max(-inf, 0) = 0) and is never stored.
template <bool WRITE_GRAD, bool ALIGNED>
__device__ __forceinline__ void stream_cls(const LevelSet &L, const StreamParams &P, const Item &it, const char *stage,
                                           int lane, float inv_tss, float &acc_cls) {
    if (ALIGNED && lane * 4 >= it.n_live) return;        // h*w % 4 == 0 => the 4 anchors of a lane are live together
    const float *tile = reinterpret_cast<const float *>(stage);
    const char *aux = stage + kTileFloats * 4;
    const int c_lo = it.group * kClsRows;
    const int c_n = min(kClsRows, P.nc - c_lo);
    // anchor j of this lane within the item
    const int a0 = ALIGNED ? lane * 4 : lane, a1 = ALIGNED ? a0 + 1 : lane + 32, a2 = ALIGNED ? a0 + 2 : lane + 64,
              a3 = ALIGNED ? a0 + 3 : lane + 96;
    const bool v0 = ALIGNED || a0 < it.n_live, v1 = ALIGNED || a1 < it.n_live, v2 = ALIGNED || a2 < it.n_live,
               v3 = ALIGNED || a3 < it.n_live;
    int4 lab4;
    float4 nr4;
    if (ALIGNED) {
        lab4 = reinterpret_cast<const int4 *>(aux)[lane];
        nr4 = reinterpret_cast<const float4 *>(aux + 512)[lane];
    } else {
        const int *la = reinterpret_cast<const int *>(aux);
        const float *na = reinterpret_cast<const float *>(aux + 512);
        lab4 = make_int4(v0 ? la[a0] : -1, v1 ? la[a1] : -1, v2 ? la[a2] : -1, v3 ? la[a3] : -1);
        nr4 = make_float4(v0 ? na[a0] : 0.f, v1 ? na[a1] : 0.f, v2 ? na[a2] : 0.f, v3 ? na[a3] : 0.f);
    }
    float *gc = WRITE_GRAD ? L.g[it.l] + ((size_t)it.b * L.C + 64 + c_lo) * it.hw + it.i0 : nullptr;
    // the class of anchor k relative to this item's first channel (never matches when background: lab = -1)
    const int r0 = lab4.x - c_lo, r1 = lab4.y - c_lo, r2 = lab4.z - c_lo, r3 = lab4.w - c_lo;
    const float gscale = P.hyp_cls * P.batch * inv_tss;
    // targets pre-multiplied by the gradient scale: grad = sig*gscale - t*gscale, and sum(x*t) = sum(x*t*gscale)/gscale
    const float inv_g = gscale != 0.f ? 1.0f / gscale : 0.f;
    const float t0 = nr4.x * gscale, t1 = nr4.y * gscale, t2 = nr4.z * gscale, t3 = nr4.w * gscale;
    // loss = sum max(x,0) - sum x*t + sum log1p(exp(-|x|)); the last sum is ONE log1p per lane of
    // S = prod(1 + e) - 1, carried as S <- S + e + S*e (four independent chains, kClsRows factors each).
    // The target is non-zero in at most one row per anchor (its label's), and only for foreground anchors: the sweep
    // treats every element as background (grad = sig*gscale) and the few label elements are patched afterwards
    // (same thread, same address, program order) -- three instructions less per element than selecting t in the sweep.
    float pos = 0.f, xt = 0.f;
    float S0 = 0.f, S1 = 0.f, S2 = 0.f, S3 = 0.f;
    auto sig_e = [&](float x, float &e) -> float {
        e = ex2_fast(fabsf(x) * -kLog2e);                   // exp(-|x|) in (0,1]
        const float inv = rcp_fast(1.0f + e);               // sigmoid(|x|)
        return x >= 0.f ? inv : e * inv;
    };
    auto elem = [&](float x, float &S) -> float {
        float e;
        const float sg = sig_e(x, e);
        S = fmaf(S, e, S + e);
        pos += fmaxf(x, 0.f);
        return sg * gscale;
    };
    auto one = [&](const float4 xv) -> float4 {
        float4 o;
        o.x = elem(xv.x, S0);
        o.y = elem(xv.y, S1);
        o.z = elem(xv.z, S2);
        o.w = elem(xv.w, S3);
        return o;
    };
    const float ninf = __int_as_float(0xff800000);
    auto row = [&](int c) {
        if (ALIGNED) {
            const float4 o = one(*reinterpret_cast<const float4 *>(tile + c * kChunk + lane * 4));
            if (WRITE_GRAD) YS_ST4(reinterpret_cast<float4 *>(gc + (size_t)c * it.hw + lane * 4), o);
        } else {
            const float *tr = tile + c * kChunk;
            const float4 o = one(make_float4(v0 ? tr[a0] : ninf, v1 ? tr[a1] : ninf, v2 ? tr[a2] : ninf, v3 ? tr[a3] : ninf));
            if (WRITE_GRAD) {
                float *gr = gc + (size_t)c * it.hw;
                if (v0) gr[a0] = o.x;
                if (v1) gr[a1] = o.y;
                if (v2) gr[a2] = o.z;
                if (v3) gr[a3] = o.w;
            }
        }
    };
    if (c_n == kClsRows) {
#pragma unroll
        for (int c = 0; c < kClsRows; ++c) row(c);
    } else {
        for (int c = 0; c < c_n; ++c) row(c);
    }
    auto patch = [&](int r, int a, float tk) {
        if (r >= 0 && r < c_n) {
            const float x = tile[r * kChunk + a];
            float e;
            const float sg = sig_e(x, e);
            xt = fmaf(x, tk, xt);
            if (WRITE_GRAD) gc[(size_t)r * it.hw + a] = fmaf(sg, gscale, -tk);
        }
    };
    patch(r0, a0, t0); patch(r1, a1, t1); patch(r2, a2, t2); patch(r3, a3, t3);
    const float S01 = fmaf(S0, S1, S0 + S1), S23 = fmaf(S2, S3, S2 + S3);
    acc_cls += pos - xt * inv_g + log1pf(fmaf(S01, S23, S01 + S23));
}

// ---- kernel ---------------------------------------------------------------------------------------------------
template <bool WRITE_GRAD, int STAGES>
__global__ void __launch_bounds__(kStreamThreads, 1) k_loss_stream(LevelSet L, StreamParams P, const __grid_constant__ StreamMaps M) {
    extern __shared__ __align__(128) char s_ring[];                 // [warp][stage][kStageBytes]
    __shared__ __align__(8) unsigned long long s_bar[kStreamWarps][kStreamStagesMax];
    __shared__ float s_sum[kStreamWarps][3];
    __shared__ double s_fin[kStreamWarps][3];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(smem_u32(&s_bar[warp][s]), 1u);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    // programmatic dependent launch: everything above ran while k_assign was still draining; its outputs are read from here on
    cudaGridDependencySynchronize();
    const float inv_tss = 1.0f / P.tss[0];
    float acc_box = 0.f, acc_cls = 0.f, acc_dfl = 0.f;
    char *ring = s_ring + (size_t)warp * STAGES * kStageBytes;
    const int step = (int)gridDim.x * kStreamWarps;
    int item = (int)blockIdx.x * kStreamWarps + warp;

    // Foreground lookup of a box item: every lane reads the assignment of "its" anchor (both half-warps the same 16).
    auto load_tg = [&](const Item &it) -> int {
        if (it.group >= 0) return -1;
#ifdef YS_EXP_NOBOX
        return -1;
#endif
        const int a16 = lane & 15;
        return a16 < it.n_live ? __ldg(P.tgi + (size_t)it.b * L.A + L.a_off[it.l] + it.i0 + a16) : -1;
    };
    // Pipeline state: cur = item being computed, nxt = item whose tile is in flight, nn = item whose tgi is in flight.
    Item cur, nxt, nn;
    int tg_cur = -1, tg_nxt = -1, tg_nn = -1;
    bool cur_loaded = false, nxt_loaded = false;
    unsigned phase_bits = 0u;                                        // bit s = parity the next wait on stage s uses
    // Schedule.  Round k covers the contiguous window of groups [k*G, (k+1)*G), G = gridDim.x, a group being kStreamWarps
    // consecutive items (warp w takes the w-th): at any time all SMs sweep ONE contiguous window of rows, which is what
    // DRAM rewards (scattering items or groups over the SMs measured 5-20 % slower).  Inside the window CTA c takes group
    // (c + k*rotate) mod G.  Without the rotation a CTA's position relative to the image it is in moves by
    // (G * kStreamWarps) mod items_per_img per round -- at the bench shape with 16 warps that is -2 items: a CTA would see
    // only class items, or only background box items, for the whole kernel and the SMs finish 20 % apart.  The host
    // picks `rotate` so that the per-round drift is ~0.38 of an image (golden-ratio sampling of the item kinds).
    const unsigned G = gridDim.x;
    const unsigned my_groups = P.rounds;
    unsigned k_nn = 0u, rot_nn = blockIdx.x % G;      // round and rotated slot of the next group to decode
    auto next_group_item = [&]() -> int {
        const unsigned g = k_nn * G + rot_nn;
        ++k_nn;
        rot_nn += P.rotate;
        if (rot_nn >= G) rot_nn -= G;
        return g < P.n_groups16 ? (int)(g * kStreamWarps) + warp : P.total_items;      // past the end: padding item
    };
    auto load_tg2 = [&](const Item &it) -> int { return it.n_live > 0 ? load_tg(it) : -1; };
    // One loaded item: (levels without a tensor map: the warp fills the tile itself first,) then the item's arithmetic.
    auto consume = [&](const Item &it, char *stage, int tg) {
        const bool tma = L.tma[it.l] != 0;
        if (!tma) fill_tile_plain(L, P, it, stage, lane);
        if (it.group >= 0) {
            if (tma) stream_cls<WRITE_GRAD, true>(L, P, it, stage, lane, inv_tss, acc_cls);
            else stream_cls<WRITE_GRAD, false>(L, P, it, stage, lane, inv_tss, acc_cls);
        } else {
            stream_box<WRITE_GRAD>(L, P, it, stage, tg, lane, inv_tss, acc_box, acc_dfl);
        }
    };
    if (STAGES == 3) {
    // Three stages: while item k is computed, the tiles of items k+1 and k+2 are in flight and the assignment lookup of
    // k+3 is on its way.  With two stages a cheap item (an all-background box item: eight stores) left the next tile
    // ~0.2 us to arrive and the warp then sat in mbar_wait (14 % of all warp samples).  The queue holds packed items.
    {
        ItemQ q0 = {0, 0, 0}, q1 = q0, q2 = q0, q3 = q0;
        int t0 = -1, t1 = -1, t2 = -1, t3 = -1;
        bool l0 = false, l1 = false;
        auto loaded_of = [&](const ItemQ &q, int tg) -> bool {
            return (q.meta & 255) > 0 && ((q.meta >> 12) > 0 || __any_sync(0xffffffffu, tg >= 0));
        };
        auto fetch = [&](ItemQ &q, int &tg) {
            const Item it = decode_item(L, P, next_group_item());
            tg = load_tg2(it);
            q = pack_item(it);
        };
        if (my_groups > 0u) {
            fetch(q0, t0);
            l0 = loaded_of(q0, t0);
            if (l0) issue_item(L, P, M, unpack_item(L, q0), ring, smem_u32(&s_bar[warp][0]), lane);
        }
        if (my_groups > 1u) {
            fetch(q1, t1);
            l1 = loaded_of(q1, t1);
            if (l1) issue_item(L, P, M, unpack_item(L, q1), ring + kStageBytes, smem_u32(&s_bar[warp][1]), lane);
        }
        if (my_groups > 2u) fetch(q2, t2);
        (void)item; (void)step; (void)cur; (void)nxt; (void)nn; (void)tg_cur; (void)tg_nxt; (void)tg_nn; (void)cur_loaded; (void)nxt_loaded;
        int s = 0;
        for (unsigned k = 0; k < my_groups; ++k) {
            const int s2 = s >= 1 ? s - 1 : 2;          // (s + 2) % 3: the stage item k-1 just left
            bool l2 = false;
            if (k + 2u < my_groups) {
                l2 = loaded_of(q2, t2);
                __syncwarp();
                fence_async_smem();
                if (l2) issue_item(L, P, M, unpack_item(L, q2), ring + (size_t)s2 * kStageBytes, smem_u32(&s_bar[warp][s2]), lane);
                if (k + 3u < my_groups) fetch(q3, t3);
            }
            const Item it = unpack_item(L, q0);
            if (l0) {
                mbar_wait(smem_u32(&s_bar[warp][s]), (phase_bits >> s) & 1u);
                phase_bits ^= 1u << s;
                consume(it, ring + (size_t)s * kStageBytes, t0);
            } else if (it.n_live > 0) {
                stream_box_zero<WRITE_GRAD>(L, it, lane);
            }
            q0 = q1; t0 = t1; l0 = l1;
            q1 = q2; t1 = t2; l1 = l2;
            q2 = q3; t2 = t3;
            s = s == 2 ? 0 : s + 1;
        }
    }
    } else {
    if (my_groups > 0u) {
        cur = decode_item(L, P, next_group_item());
        tg_cur = load_tg2(cur);
        cur_loaded = cur.n_live > 0 && (cur.group >= 0 || __any_sync(0xffffffffu, tg_cur >= 0));
        if (cur_loaded) issue_item(L, P, M, cur, ring, smem_u32(&s_bar[warp][0]), lane);
    }
    if (my_groups > 1u) {
        nxt = decode_item(L, P, next_group_item());
        tg_nxt = load_tg2(nxt);
    }
    (void)item; (void)step;
    for (unsigned k = 0; k < my_groups; ++k) {
        const int s = (int)(k & 1u);
        // ---- launch the next item's tile into the other stage (its previous reader finished last iteration)
        const bool has_nxt = k + 1u < my_groups;
        if (has_nxt) {
            nxt_loaded = nxt.n_live > 0 && (nxt.group >= 0 || __any_sync(0xffffffffu, tg_nxt >= 0));
            __syncwarp();
            fence_async_smem();
            if (nxt_loaded) issue_item(L, P, M, nxt, ring + (size_t)(s ^ 1) * kStageBytes, smem_u32(&s_bar[warp][s ^ 1]), lane);
            // and the assignment lookup of the one after it
            if (k + 2u < my_groups) {
                nn = decode_item(L, P, next_group_item());
                tg_nn = load_tg2(nn);
            }
        }
        // ---- compute the current item
        if (cur_loaded) {
            mbar_wait(smem_u32(&s_bar[warp][s]), (phase_bits >> s) & 1u);
            phase_bits ^= 1u << s;
            consume(cur, ring + (size_t)s * kStageBytes, tg_cur);
        } else if (cur.n_live > 0) {
            stream_box_zero<WRITE_GRAD>(L, cur, lane);
        }
        cur = nxt; tg_cur = tg_nxt; cur_loaded = nxt_loaded;
        nxt = nn; tg_nxt = tg_nn;
    }

    }

    // ------------------------------------------------------------------ CTA partial sums (fixed order)
    acc_box = warp_sum(acc_box);
    acc_cls = warp_sum(acc_cls);
    acc_dfl = warp_sum(acc_dfl);
    if (lane == 0) { s_sum[warp][0] = acc_box; s_sum[warp][1] = acc_cls; s_sum[warp][2] = acc_dfl; }
    __syncthreads();
    if (tid == 0) {
        float t0 = 0.f, t1 = 0.f, t2 = 0.f;
#pragma unroll
        for (int w = 0; w < kStreamWarps; ++w) { t0 += s_sum[w][0]; t1 += s_sum[w][1]; t2 += s_sum[w][2]; }
        P.partials[blockIdx.x] = make_float4(t0, t1, t2, 0.f);
        __threadfence();
        const unsigned prev = atomicAdd(P.done, 1u);
        s_last = (prev == gridDim.x - 1u);
    }
    __syncthreads();
    if (!s_last) return;
    // ------------------------------------------------------------------ last CTA: ordered fp64 reduction
    __threadfence();
    double a0 = 0, a1 = 0, a2 = 0;
    for (int i = tid; i < (int)gridDim.x; i += kStreamThreads) {
        const float4 v = __ldcg(P.partials + i);
        a0 += v.x; a1 += v.y; a2 += v.z;
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2);
    if (lane == 0) { s_fin[warp][0] = a0; s_fin[warp][1] = a1; s_fin[warp][2] = a2; }
    __syncthreads();
    if (tid == 0) {
        double t0 = 0, t1 = 0, t2 = 0;
        for (int w = 0; w < kStreamWarps; ++w) { t0 += s_fin[w][0]; t1 += s_fin[w][1]; t2 += s_fin[w][2]; }
        const double inv = 1.0 / (double)P.tss[0];
        // loss[0] = box/tss*hyp_box, loss[1] = cls/tss*hyp_cls, loss[2] = dfl/tss*hyp_dfl (:475, :300, :306, :495-497)
        const float lb = (float)(t0 * inv) * P.hyp_box;
        const float lc = (float)(t1 * inv) * P.hyp_cls;
        const float ld = (float)(t2 * inv) * P.hyp_dfl;
        P.loss_out[0] = (lb + lc + ld) * P.batch;   // loss.sum() * batch_size (:499)
        P.loss_out[1] = lb;
        P.loss_out[2] = lc;
        P.loss_out[3] = ld;
        *P.done = 0u;
    }
}

}  // namespace ys
