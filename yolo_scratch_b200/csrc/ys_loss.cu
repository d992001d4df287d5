// ys_loss.cu -- fused detection loss (assign + BCE + CIoU + DFL, forward and backward) and the dense
// TaskAlignedAssigner entry point.  Replaces losses/yolo_v8_loss.py of the reference; see include/ys_b200.h.
//
// Launch sequence of ys_detection_loss (one stream, no host sync, three kernels):
//   k_pack_gt   flat GT list -> per-image CSR (stable), pixel boxes                     [replaces preprocess :501-527]
//   k_assign    decode + in-box / IoU scan + preliminary assignment; the last block of each image replays the
//               reference's forced-assignment loops; the last image computes target_scores_sum  [reads 64 box channels]
//   k_loss      BCE + CIoU + DFL forward and backward, NCHW in, NCHW out; the last block reduces the partial
//               sums in a fixed order                                                   [reads 144 ch, writes 144 ch]
#include <string.h>
#include "ys_assign.cuh"
#include "ys_loss_stream.cuh"

namespace ys {

// =====================================================================================================
// GT packing
// =====================================================================================================
constexpr int kPackThreads = 1024;
constexpr int kPackSmemB = 8192;
#ifndef YS_RISK_MAX_CENTRES
#define YS_RISK_MAX_CENTRES 32
#endif
constexpr int kRiskMaxCentres = YS_RISK_MAX_CENTRES;
constexpr int kContainMaxGt = 64;     // images with at most this many GT get the containment hint
constexpr int kOrderMaxB = 1024;      // batches up to this size get their images ordered "fix-up likely" first   // GT with at most this many anchor centres inside get their best-IoU anchor from the scan

// Number of grid centres c_i = (i + 0.5) * stride, i in [0, n), inside (lo, hi): an ESTIMATE (a centre sitting on an
// edge may be counted either way).  It only feeds k_assign's launch-order hint, never a result, and k_pack_gt is one
// block that runs once per step with a cold instruction cache -- the exact edge walk that used to be here cost 5 us
// of the step's critical path.
__device__ __forceinline__ int centres_inside(float lo, float hi, int n, float stride) {
    const float inv = __frcp_rn(stride);
    const int i0 = max((int)floorf(lo * inv - 0.5f) + 1, 0);
    const int i1 = min((int)ceilf(hi * inv - 0.5f) - 1, n - 1);
    return max(0, i1 - i0 + 1);
}

// The general packing, one block of kPackThreads threads: any row order, rows with an image id outside [0, B) ignored,
// any N and B.  Shared arrays are the caller's (s_cursor[kPackSmemB], s_weight[kOrderMaxB], s_part[32], s_unsorted[1]).
__device__ void pack_gt_general(const LevelSet &L, const long long *__restrict__ batch_idx, const long long *__restrict__ cls,
                                const float4 *__restrict__ bboxes, int N, int B, int nc, float sx, float sy,
                                int *__restrict__ gt_off, int *__restrict__ rank, int *__restrict__ cursor_g,
                                float4 *__restrict__ gt_box, int *__restrict__ gt_label, unsigned char *__restrict__ risk,
                                unsigned long long *__restrict__ best_key, int *__restrict__ owned_cnt,
                                int *__restrict__ gt_flags, unsigned *__restrict__ counters, int n_counters,
                                int *__restrict__ img_order, int n_identity, int *s_cursor, int *s_weight, int *s_part,
                                int *s_unsorted_p) {
    int &s_unsorted = *s_unsorted_p;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    volatile int *cursor = (B <= kPackSmemB) ? (volatile int *)s_cursor : (volatile int *)cursor_g;
    for (int i = tid; i < B; i += kPackThreads) cursor[i] = 0;
    for (int i = tid; i < n_counters; i += kPackThreads) counters[i] = 0u;    // arrival counters of k_assign / k_loss
    for (int i = tid; i < kOrderMaxB; i += kPackThreads) s_weight[i] = 0;
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
        best_key[i] = 0xFFFFFFFFull;      // = argmax_key(IoU 0, anchor 0): the first maximum of an all-zero column
        owned_cnt[i] = 0;
        gt_flags[i] = 0;
        long long id = batch_idx[i];
        if (id < 0 || id >= B) continue;
        int dst = sorted ? i : (gt_off[(int)id] + rank[i]);
        float4 q = bboxes[i];
        // targets_bbox * [sx, sy, sx, sy]   (losses/yolo_v8_loss.py:515-516)
        const float4 qp = make_float4(YS_MUL(q.x, sx), YS_MUL(q.y, sy), YS_MUL(q.z, sx), YS_MUL(q.w, sy));
        gt_box[dst] = qp;
        // hint for k_assign: GT with few anchor centres inside are the ones the forced-assignment loops
        // (:117-139, :146-162) can fire for; the scan computes their best-IoU anchor on the way.
        int centres = 0;
        for (int l = 0; l < L.n && centres <= kRiskMaxCentres; ++l)
            centres += centres_inside(qp.x, qp.z, L.w[l], L.stride[l]) * centres_inside(qp.y, qp.w, (int)((float)L.hw[l] * __frcp_rn((float)L.w[l]) + 0.5f), L.stride[l]);
        risk[dst] = centres <= kRiskMaxCentres ? 1 : 0;
        if (B <= kOrderMaxB) atomicAdd(&s_weight[(int)id], (centres <= kRiskMaxCentres ? (centres <= 4 ? 64 : 1) : 0) * 2048 + 1);
        long long c = cls[i];
        c = c < 0 ? 0 : (c > nc - 1 ? nc - 1 : c);
        gt_label[dst] = (int)c;
    }
    // Second "at risk" hint (small images only): GT j lies inside another GT's box -- every anchor centre inside j is
    // then inside that GT too, and j owns nothing if it loses all those arg-max contests (the loop-2 case, :146-162).
    __syncthreads();
    for (int i = tid; i < N; i += kPackThreads) {
        const long long id = batch_idx[i];
        if (id < 0 || id >= B) continue;
        const int off = gt_off[(int)id], cnt = gt_off[(int)id + 1] - off;
        const int dst = sorted ? i : (off + rank[i]);
        if (cnt > kContainMaxGt || risk[dst]) continue;
        const float4 q = gt_box[dst];
        bool inside = false;
        for (int k = off; k < off + cnt; ++k) {
            const float4 o = gt_box[k];
            inside |= (k != dst) && o.x <= q.x && o.y <= q.y && o.z >= q.z && o.w >= q.w;
        }
        if (inside) {
            risk[dst] = 1;
            if (B <= kOrderMaxB) atomicAdd(&s_weight[(int)id], 2048);
        }
    }
    // Launch order of the images in k_assign (blockIdx.y -> image): the ones whose forced-assignment fix-up is
    // likely to have work first, so that those fix-ups run under the scan of the remaining images instead of
    // after it.  A scheduling hint only: results do not depend on the order.
    __syncthreads();
    for (int b = tid; b < B; b += kPackThreads) {
        int r = b;       // the first n_identity slots keep their own image (k_assign decodes those before it may read this)
        if (B <= kOrderMaxB && b >= n_identity) {
            const int wb = s_weight[b];
            r = n_identity;
            for (int o = n_identity; o < B; ++o) {
                const int wo = s_weight[o];
                r += (wo > wb || (wo == wb && o < b)) ? 1 : 0;
            }
        }
        img_order[r] = b;
    }
}

// ---------------------------------------------------------------------------------------------------
// GT packing above 1024 GT or 1024 images (the dense 1280 px scene of BASELINE configs[4]: 19200 GT per batch).
// Multi-block, one row per thread.  What the reference collate produces -- rows sorted by image, every id valid
// (datasets/crater_dataset_cuda.py:327-341) -- needs no coordination between blocks: the packed position of row i is i,
// gt_off[b] is the first row of image b (written by the thread that sees the boundary), the hints come from the
// row's own box and from its image's other rows, which lie within 64 rows of it whenever the containment hint applies
// (read from a halo'd shared-memory copy).  Every block checks its rows' order; the last block to finish (arrival
// counter, reset by that block) either ranks the images for k_assign's launch order or, if any block saw a row out of
// order or an invalid id, redoes the whole packing with pack_gt_general (one block; r02's single-block kernel took 70 us
// of the 1.03 ms configs[4] step this way).  Same outputs as pack_gt_general, bit for bit.  pack_ctr = {arrivals, redo
// flag, launch weight of image 0, 1, ...}: zeroed by a memset node in front of this kernel.
// ---------------------------------------------------------------------------------------------------
constexpr int kPackHalo = kContainMaxGt;                       // rows of the neighbouring blocks each block also loads
constexpr int kPackRaw = (kPackSmemB + kOrderMaxB + 32 + 4) * 4;   // bytes: the general path's arrays (>= the sorted path's)
static_assert((kPackThreads + 2 * kPackHalo) * (4 + 16) <= kPackRaw, "sorted-path staging must fit the shared buffer");

__global__ void __launch_bounds__(kPackThreads)
k_pack_gt(LevelSet L, const long long *__restrict__ batch_idx, const long long *__restrict__ cls, const float4 *__restrict__ bboxes,
          int N, int B, int nc, float sx, float sy, int *__restrict__ gt_off, int *__restrict__ rank,
          int *__restrict__ cursor_g, float4 *__restrict__ gt_box, int *__restrict__ gt_label,
          unsigned char *__restrict__ risk, unsigned long long *__restrict__ best_key, int *__restrict__ owned_cnt,
          int *__restrict__ gt_flags, unsigned *__restrict__ counters, int n_counters, int *__restrict__ img_order,
          int n_identity, unsigned *__restrict__ pack_ctr) {
    int *weight_g = reinterpret_cast<int *>(pack_ctr + 2);
    __shared__ __align__(16) unsigned char s_raw[kPackRaw];
    __shared__ int s_last;
    const int tid = threadIdx.x;
    cudaTriggerProgrammaticLaunchCompletion();      // k_assign may start its prologue (it waits for this grid to finish)
    // ---- sorted-input pass: rows [row0 - halo, row0 + 1024 + halo) staged in shared memory
    float4 *s_box = reinterpret_cast<float4 *>(s_raw);                              // [1024 + 2 halo]
    int *s_id = reinterpret_cast<int *>(s_raw + (kPackThreads + 2 * kPackHalo) * 16);   // [1024 + 2 halo], -1 outside [0, N)
    const int row0 = blockIdx.x * kPackThreads;
    const bool order_images = B <= kOrderMaxB;
    for (int k = tid; k < kPackThreads + 2 * kPackHalo; k += kPackThreads) {
        const int r = row0 - kPackHalo + k;
        int id = -1;
        float4 qp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r >= 0 && r < N) {
            const long long id64 = batch_idx[r];
            id = (id64 >= 0 && id64 < B) ? (int)id64 : -2;
            const float4 q = bboxes[r];
            qp = make_float4(YS_MUL(q.x, sx), YS_MUL(q.y, sy), YS_MUL(q.z, sx), YS_MUL(q.w, sy));   // (losses/yolo_v8_loss.py:515-516)
        }
        s_id[k] = id;
        s_box[k] = qp;
    }
    __syncthreads();
    const int i = row0 + tid, k = tid + kPackHalo;
    if (i < N) {
        const int id = s_id[k];
        const int prev = i > 0 ? s_id[k - 1] : -1;             // image of row i-1 (-1: there is none, -2: invalid id)
        const bool bad = id < 0 || (i > 0 && (prev < 0 || prev > id));
        if (bad) {
            pack_ctr[1] = 1u;                                  // someone has to redo this properly
        } else {
            for (int j = prev + 1; j <= id; ++j) gt_off[j] = i;        // images prev+1 .. id start at row i
            if (i == N - 1)
                for (int j = id + 1; j <= B; ++j) gt_off[j] = N;
            best_key[i] = 0xFFFFFFFFull;      // = argmax_key(IoU 0, anchor 0): the first maximum of an all-zero column
            owned_cnt[i] = 0;
            gt_flags[i] = 0;
            const float4 qp = s_box[k];
            gt_box[i] = qp;
            long long c = cls[i];
            c = c < 0 ? 0 : (c > nc - 1 ? nc - 1 : c);
            gt_label[i] = (int)c;
            int centres = 0;
            for (int l = 0; l < L.n && centres <= kRiskMaxCentres; ++l)
                centres += centres_inside(qp.x, qp.z, L.w[l], L.stride[l]) * centres_inside(qp.y, qp.w, (int)((float)L.hw[l] * __frcp_rn((float)L.w[l]) + 0.5f), L.stride[l]);
            int rk = centres <= kRiskMaxCentres ? 1 : 0;
            int wgt = (centres <= kRiskMaxCentres ? (centres <= 4 ? 64 : 1) : 0) * 2048 + 1;
            if (!rk) {
                // containment hint (images with at most kContainMaxGt GT): the image's rows are within the halo
                int lo = k, hi = k + 1;
                while (lo > k - kPackHalo && s_id[lo - 1] == id) --lo;
                while (hi < k + kPackHalo + 1 && s_id[hi] == id) ++hi;
                if (hi - lo <= kContainMaxGt) {            // more rows than that: one of the two walks ran its full length
                    bool inside = false;
                    for (int m = lo; m < hi; ++m) {
                        const float4 o = s_box[m];
                        inside |= (m != k) && o.x <= qp.x && o.y <= qp.y && o.z >= qp.z && o.w >= qp.w;
                    }
                    if (inside) { rk = 1; wgt += 2048; }
                }
            }
            risk[i] = (unsigned char)rk;
            if (order_images) atomicAdd(&weight_g[id], wgt);
        }
    }
    if (N == 0 && blockIdx.x == 0)
        for (int j = tid; j <= B; j += kPackThreads) gt_off[j] = 0;
    // ---- arrival; the last block finishes the job
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(&pack_ctr[0], 1u) == gridDim.x - 1) ? 1 : 0;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const bool redo = *reinterpret_cast<volatile unsigned *>(&pack_ctr[1]) != 0u;
    int *s_w = reinterpret_cast<int *>(s_raw);
    if (!redo && order_images)
        for (int b = tid; b < B; b += kPackThreads) s_w[b] = __ldcg(&weight_g[b]);
    __syncthreads();
    if (redo) {
        int *s_cursor = reinterpret_cast<int *>(s_raw);
        int *s_weight = s_cursor + kPackSmemB;
        int *s_part = s_weight + kOrderMaxB;
        pack_gt_general(L, batch_idx, cls, bboxes, N, B, nc, sx, sy, gt_off, rank, cursor_g, gt_box, gt_label, risk, best_key,
                        owned_cnt, gt_flags, counters, n_counters, img_order, n_identity, s_cursor, s_weight, s_part,
                        s_part + 32);
        return;
    }
    for (int c = tid; c < n_counters; c += kPackThreads) counters[c] = 0u;    // arrival counters of k_assign / k_loss
    // launch order of the images in k_assign: heaviest expected fix-up first (a scheduling hint only)
    for (int b = tid; b < B; b += kPackThreads) {
        int r = b;       // the first n_identity slots keep their own image (k_assign decodes those before it may read this)
        if (order_images && b >= n_identity) {
            const int wb = s_w[b];
            r = n_identity;
            for (int o = n_identity; o < B; ++o) {
                const int wo = s_w[o];
                r += (wo > wb || (wo == wb && o < b)) ? 1 : 0;
            }
        }
        img_order[r] = b;
    }
}

// ---------------------------------------------------------------------------------------------------
// Fast path of the GT packing: at most 1024 GT and 1024 images (every training batch of the bench shapes).
// One GT per thread, one global round trip: the list is read into registers once, everything in between -- per-image
// counts, stable ranks, the exclusive scan, risk hints, containment, launch order -- lives in shared memory.
// Same outputs as k_pack_gt, bit for bit.
// ---------------------------------------------------------------------------------------------------
constexpr int kSmallMax = 1024;
__global__ void __launch_bounds__(kSmallMax)
k_pack_gt_small(LevelSet L, const long long *__restrict__ batch_idx, const long long *__restrict__ cls,
                const float4 *__restrict__ bboxes, int N, int B, int nc, float sx, float sy, int *__restrict__ gt_off,
                float4 *__restrict__ gt_box, int *__restrict__ gt_label, unsigned char *__restrict__ risk,
                unsigned long long *__restrict__ best_key, int *__restrict__ owned_cnt, int *__restrict__ gt_flags,
                unsigned *__restrict__ counters, int n_counters, int *__restrict__ img_order, int n_identity) {
    __shared__ int s_id[kSmallMax];          // image of source row i (-1: ignored)
    __shared__ int s_cnt[kSmallMax];         // GT per image, then reused as launch weights
    __shared__ int s_off[kSmallMax + 1];     // exclusive scan of the counts
    __shared__ float4 s_box[kSmallMax];      // pixel boxes in destination order
    __shared__ unsigned char s_risk[kSmallMax];
    __shared__ int s_part[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    cudaTriggerProgrammaticLaunchCompletion();      // k_assign may start its prologue (it waits for this grid to finish)
    // ---- the only global reads
    long long id64 = -1, c64 = 0;
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid < N) {
        id64 = batch_idx[tid];
        c64 = cls[tid];
        q = bboxes[tid];
    }
    const bool ok = tid < N && id64 >= 0 && id64 < B;
    const int id = ok ? (int)id64 : -1;
    s_id[tid] = id;
    s_cnt[tid] = 0;
    for (int i = tid; i < n_counters; i += kSmallMax) counters[i] = 0u;    // arrival counters of k_assign / k_loss
    __syncthreads();
    // ---- stable rank within the image.  Sorted input (what the reference collate produces): rank = i - first row of
    //      the image; otherwise count the earlier rows of the same image.
    const bool in_order = !(tid > 0 && tid < N && (s_id[tid - 1] > id || s_id[tid - 1] < 0)) && (tid >= N || ok);
    const int sorted = __syncthreads_and(in_order);
    int rank = 0;
    if (ok) {
        if (sorted) {
            int first = tid;            // walk back to the first row of this image: bounded by the image's GT count
            while (first > 0 && s_id[first - 1] == id) --first;
            rank = tid - first;
        } else {
            for (int j = 0; j < tid; ++j) rank += (s_id[j] == id) ? 1 : 0;
        }
        atomicAdd(&s_cnt[id], 1);
    }
    __syncthreads();
    // ---- exclusive scan of the per-image counts (one image per thread)
    const int cnt_b = tid < B ? s_cnt[tid] : 0;
    int incl = cnt_b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_part[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const int v = s_part[lane];
        int inc2 = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc2, o);
            if (lane >= o) inc2 += t;
        }
        s_part[lane] = inc2 - v;
    }
    __syncthreads();
    const int excl = s_part[warp] + incl - cnt_b;
    if (tid < B) { s_off[tid] = excl; gt_off[tid] = excl; }
    if (tid == B - 1) { s_off[B] = excl + cnt_b; gt_off[B] = excl + cnt_b; }
    if (tid < N) { best_key[tid] = 0xFFFFFFFFull; owned_cnt[tid] = 0; gt_flags[tid] = 0; }   // key (IoU 0, anchor 0)
    __syncthreads();
    // ---- placement, pixel scale (losses/yolo_v8_loss.py:515-516), risk hint from the anchor grid
    int dst = -1;
    if (ok) {
        dst = s_off[id] + rank;
        const float4 qp = make_float4(YS_MUL(q.x, sx), YS_MUL(q.y, sy), YS_MUL(q.z, sx), YS_MUL(q.w, sy));
        gt_box[dst] = qp;
        s_box[dst] = qp;
        long long c = c64 < 0 ? 0 : (c64 > nc - 1 ? nc - 1 : c64);
        gt_label[dst] = (int)c;
        int centres = 0;
        for (int l = 0; l < L.n && centres <= kRiskMaxCentres; ++l)
            centres += centres_inside(qp.x, qp.z, L.w[l], L.stride[l]) * centres_inside(qp.y, qp.w, (int)((float)L.hw[l] * __frcp_rn((float)L.w[l]) + 0.5f), L.stride[l]);
        s_risk[dst] = centres <= kRiskMaxCentres ? (centres <= 4 ? 2 : 1) : 0;
    }
    if (tid < B) s_cnt[tid] = 0;            // from here on: launch weights
    __syncthreads();
    // ---- containment hint + launch weights
    if (ok) {
        const int off = s_off[id], cnt = s_off[id + 1] - off;
        int rk = s_risk[dst];
        int wgt = rk == 2 ? 64 : rk;
        if (!rk && cnt <= kContainMaxGt) {
            const float4 me = s_box[dst];
            bool inside = false;
            for (int k = off; k < off + cnt; ++k) {
                const float4 o = s_box[k];
                inside |= (k != dst) && o.x <= me.x && o.y <= me.y && o.z >= me.z && o.w >= me.w;
            }
            if (inside) { rk = 1; wgt = 1; }
        }
        risk[dst] = rk ? 1 : 0;
        // secondary key: the image's GT count (more GT, more arg-max contests lost: ownerless GT are what the hint
        // cannot see), so the images launched last are the ones with the fewest GT
        atomicAdd(&s_cnt[id], wgt * 2048 + 1);
    }
    __syncthreads();
    // ---- launch order of the images in k_assign: heaviest expected fix-up first (a scheduling hint only)
    if (tid < B) {
        int r = tid;     // the first n_identity slots keep their own image (k_assign decodes those before it may read this)
        if (tid >= n_identity) {
            const int wb = s_cnt[tid];
            r = n_identity;
            for (int o = n_identity; o < B; ++o) {
                const int wo = s_cnt[o];
                r += (wo > wb || (wo == wb && o < tid)) ? 1 : 0;
            }
        }
        img_order[r] = tid;
    }
}

__global__ void k_scale_inplace(float *__restrict__ data, size_t n, const float *__restrict__ scale) {
    const float s = *scale;
    if (s == 1.0f) return;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) data[i] *= s;
}

// Per-anchor state from the padded anchor space (workspace) to the reference's dense numbering (caller's buffers).
__global__ void k_export_state(LevelSet L, int B, const float4 *__restrict__ pbox, const int *__restrict__ tgi,
                               const float *__restrict__ norm, float4 *__restrict__ out_box, int *__restrict__ out_tgi,
                               float *__restrict__ out_norm) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)B * L.A_true) return;
    const int b = (int)(i / L.A_true), a = (int)(i - (size_t)b * L.A_true);
    int l = 0;
#pragma unroll
    for (int k = 1; k < YS_MAX_LEVELS; ++k)
        if (k < L.n && a >= L.a_true_off[k]) l = k;
    const size_t src = (size_t)b * L.A + L.a_off[l] + (a - L.a_true_off[l]);
    if (out_box) out_box[i] = pbox[src];
    if (out_tgi) out_tgi[i] = tgi[src];
    if (out_norm) out_norm[i] = norm[src];
}

// Dense assigner outputs (losses/yolo_v8_loss.py:246-270): labels / boxes for ALL anchors (background gets
// GT slot 0), one-hot * norm scores, bool fg, i64 gt index.
__global__ void k_assign_emit(int B, int A, int G, int nc, const int *__restrict__ tgi, const float *__restrict__ norm,
                              const float *__restrict__ gt_labels, const float4 *__restrict__ gt_bboxes,
                              float *__restrict__ target_labels, float4 *__restrict__ target_bboxes,
                              float *__restrict__ target_scores, unsigned char *__restrict__ fg_mask,
                              long long *__restrict__ target_gt_idx) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < (size_t)B * A;                               // no early return: the score rows are a warp's joint work
    int g = -1;
    float lab = 0.f;
    if (live) {
        const int b = (int)(i / A);
        g = tgi[i];
        const int slot = g < 0 ? 0 : g;
        lab = gt_labels[(size_t)b * G + slot];
        lab = lab < 0.f ? 0.f : (lab > (float)nc ? (float)nc : lab);   // clamp_(0, num_classes) (:256)
        target_labels[i] = lab;
        target_bboxes[i] = gt_bboxes[(size_t)b * G + slot];
        fg_mask[i] = g >= 0;
        target_gt_idx[i] = slot;
    }
    if (target_scores && i - (threadIdx.x & 31) < (size_t)B * A) {     // warp-uniform
        // the warp's 32 rows of nc floats are one contiguous run: written by all lanes together (coalesced; float4 when the
        // run is 16-byte aligned), each element's (label, value) fetched from the lane that owns its anchor.  One thread per
        // row (32 addresses 4 * nc bytes apart per store) ran the entry point at 0.35 ms for B=64.
        const int li = (int)lab;
        const float v = (live && g >= 0) ? norm[i] : 0.f;
        const int lane = threadIdx.x & 31;
        const size_t i0 = i - lane;                                   // first anchor of the warp (blockDim.x % 32 == 0)
        const int rows = (int)min((size_t)32, (size_t)B * A - i0);
        float *dst = target_scores + i0 * nc;
        const int total = rows * nc;
        if ((nc & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            for (int base = 0; base < total; base += 128) {            // warp-uniform trip count: the shuffles need all lanes
                const int e = base + lane * 4, ec = min(e, total - 4);
                const int r = ec / nc, c = ec - r * nc;                // four consecutive classes of one anchor (nc % 4 == 0)
                const int lr = __shfl_sync(0xffffffffu, li, r);
                const float vr = __shfl_sync(0xffffffffu, v, r);
                if (e < total)
                    *reinterpret_cast<float4 *>(dst + e) = make_float4(c == lr ? vr : 0.f, c + 1 == lr ? vr : 0.f,
                                                                       c + 2 == lr ? vr : 0.f, c + 3 == lr ? vr : 0.f);
            }
        } else {
            for (int base = 0; base < total; base += 32) {
                const int e = base + lane, ec = min(e, total - 1);
                const int r = ec / nc, c = ec - r * nc;
                const int lr = __shfl_sync(0xffffffffu, li, r);
                const float vr = __shfl_sync(0xffffffffu, v, r);
                if (e < total) dst[e] = c == lr ? vr : 0.f;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------
// Top-k mode of the dense assigner (non-default; north_star "alignment-metric top-k", SURVEY D1 -- the reference stores
// `topk` and never reads it, losses/yolo_v8_loss.py:70-76): per GT, keep only the k anchors with the largest alignment
// metric score[label]^alpha * overlap^beta among the anchors whose centre lies inside it (metric > 0).
// One block per (GT, image).  The k-th largest metric is found by a 4-pass radix select on the metric's bit pattern
// (positive floats order as unsigned integers; 256-bin histogram per pass, the metric is recomputed instead of stored),
// then one ordered pass writes the selection bitmask row (one ballot word per warp).  Ties at the k-th value go to the
// lower anchor index (torch.topk leaves them unspecified).
// ---------------------------------------------------------------------------------------------------
constexpr int kTopkThreads = 256;
__global__ void __launch_bounds__(kTopkThreads)
k_topk_select(int A, int G, int nc, const float *__restrict__ pd_scores, const float4 *__restrict__ pd_box,
              const float2 *__restrict__ anc, const float *__restrict__ gt_labels, const float4 *__restrict__ gt_box,
              const float *__restrict__ mask_gt, float alpha, float beta, int ciou, int topk, unsigned *__restrict__ sel,
              int sel_words) {
    __shared__ int s_hist[256];
    __shared__ unsigned s_prefix;
    __shared__ int s_need, s_eq, s_all;
    const int g = blockIdx.x, b = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned *out = sel + ((size_t)b * G + g) * sel_words;
    if (mask_gt[(size_t)b * G + g] == 0.f) {
        for (int w = tid; w < sel_words; w += kTopkThreads) out[w] = 0u;
        return;
    }
    const float4 q = gt_box[(size_t)b * G + g];
    const float area_q = area_eps_h(q.x, q.y, q.z, q.w);
    int lab = (int)gt_labels[(size_t)b * G + g];
    lab = lab < 0 ? 0 : (lab > nc - 1 ? nc - 1 : lab);
    AssignIO io_m;                     // only the metric switch of pair_ov is read
    io_m.ciou = ciou;
    auto metric_bits = [&](int a) -> unsigned {
        if (a >= A) return 0u;
        const float2 c = anc[a];
        if (!centre_in_box(c.x, c.y, q)) return 0u;
        const float4 p = pd_box[(size_t)b * A + a];
        const float ov = pair_ov<false, false>(io_m, p, area_eps_h(p.x, p.y, p.z, p.w), q, area_q);
        const float sc = pd_scores[((size_t)b * A + a) * nc + lab];
        const float m = pow_special(sc, alpha) * pow_special(ov, beta);      // align_metric (:206)
        return m > 0.f ? __float_as_uint(m) : 0u;
    };
    if (tid == 0) { s_prefix = 0u; s_need = topk; s_all = 0; s_eq = 0; }
    for (int pass = 0; pass < 4; ++pass) {
        for (int i = tid; i < 256; i += kTopkThreads) s_hist[i] = 0;
        __syncthreads();
        if (s_all) break;                                   // fewer positive candidates than k: everything is selected
        const int shift = 24 - 8 * pass;
        const unsigned prefix = s_prefix;
        for (int a = tid; a < A; a += kTopkThreads) {
            const unsigned bits = metric_bits(a);
            if (bits != 0u && (pass == 0 || (bits >> (shift + 8)) == prefix)) atomicAdd(&s_hist[(bits >> shift) & 255u], 1);
        }
        __syncthreads();
        if (tid == 0) {
            int need = s_need, bin = 255, cum = 0;
            for (; bin >= 0; --bin) {
                if (cum + s_hist[bin] >= need) break;
                cum += s_hist[bin];
            }
            if (bin < 0) {
                s_all = 1;                                  // only possible in pass 0
            } else {
                s_need = need - cum;
                s_eq = s_hist[bin];
                s_prefix = (prefix << 8) | (unsigned)bin;
            }
        }
        __syncthreads();
    }
    const bool all = s_all != 0;
    const unsigned T = all ? 1u : s_prefix;                 // the k-th largest metric (bit pattern)
    const bool take_eq = all || s_eq <= s_need;             // every candidate equal to T fits
    for (int base = 0; base < sel_words * 32; base += kTopkThreads) {
        const int a = base + tid;
        const unsigned bits = metric_bits(a);
        const bool on = bits != 0u && (bits > T || (bits == T && take_eq));
        const unsigned m = __ballot_sync(0xffffffffu, on);
        const int w = (base >> 5) + warp;
        if (lane == 0 && w < sel_words) out[w] = m;
    }
    if (!take_eq) {                                          // ties at the k-th value: the lowest anchors first
        __syncthreads();
        if (tid == 0) {
            int need = s_need;
            for (int a = 0; a < A && need > 0; ++a)
                if (metric_bits(a) == T) { out[a >> 5] |= 1u << (a & 31); --need; }
        }
    }
}

static int build_levels(LevelSet &L, int n_levels, const float *const *xs, float *const *gs, const int *hs,
                        const int *ws, const float *strides, int nc, int reg_max) {
    YS_REQUIRE(n_levels >= 1 && n_levels <= YS_MAX_LEVELS, YS_ERR_BAD_ARG, "n_levels must be in [1,%d]", YS_MAX_LEVELS);
    YS_REQUIRE(reg_max == 16, YS_ERR_UNSUPPORTED, "reg_max must be 16 (reference Detect fixes it, models/yolo11_modules.py:214)");
    YS_REQUIRE(nc >= 1, YS_ERR_BAD_ARG, "nc must be >= 1");
    L.n = n_levels;
    L.C = 4 * reg_max + nc;
    int a = 0, at = 0, ch = 0, hs16 = 0, sb = 0;
    for (int i = 0; i < n_levels; ++i) {
        YS_REQUIRE(xs[i] != nullptr && hs[i] > 0 && ws[i] > 0, YS_ERR_BAD_ARG, "level %d: null pointer or empty shape", i);
        YS_REQUIRE(((uintptr_t)xs[i] & 3) == 0, YS_ERR_BAD_ARG, "level %d: pointer must be 4-byte aligned", i);
        if (gs) YS_REQUIRE(gs[i] != nullptr && ((uintptr_t)gs[i] & 3) == 0, YS_ERR_BAD_ARG, "level %d: bad grad pointer", i);
        const int hw = hs[i] * ws[i];
        // Any h*w is accepted (the reference only views and concatenates, losses/yolo_v8_loss.py:391-396).  Levels whose
        // rows are 16-byte aligned take the TMA loaders and float4 stores; the others the plain-load loaders.
        L.tma[i] = (hw % 4 == 0) && (((uintptr_t)xs[i] & 15) == 0) && (!gs || ((uintptr_t)gs[i] & 15) == 0);
#ifdef YS_FORCE_NO_TMA       // test builds: every level through the plain-load loaders
        L.tma[i] = 0;
#endif
        L.hseg_off[i] = hs16;
        L.sblk_off[i] = sb;
        sb += (hw + kScanThreads - 1) / kScanThreads;
        hs16 += (hw + kBoxItem - 1) / kBoxItem;
        L.x[i] = xs[i];
        L.g[i] = gs ? gs[i] : nullptr;
        L.w[i] = ws[i];
        L.hw[i] = hw;
        L.stride[i] = strides[i];
        L.a_off[i] = a;
        L.a_true_off[i] = at;
        L.chunk_off[i] = ch;
        a += (hw + 3) & ~3;
        at += hw;
        ch += (hw + kChunk - 1) / kChunk;
    }
    for (int i = n_levels; i <= YS_MAX_LEVELS; ++i) { L.a_off[i] = a; L.a_true_off[i] = at; L.chunk_off[i] = ch; L.hseg_off[i] = hs16; L.sblk_off[i] = sb; }
    for (int i = n_levels; i < YS_MAX_LEVELS; ++i) { L.x[i] = nullptr; L.g[i] = nullptr; L.w[i] = 1; L.hw[i] = 0; L.stride[i] = 1.f; L.tma[i] = 0; }
    L.A = a;
    L.A_true = at;
    L.n_chunks = ch;
    L.n_hsegs = hs16;
    L.n_sblk = sb;
    return YS_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// Launch with programmatic dependent launch (PDL): the kernel may start while its predecessor in the stream is still
// draining (the predecessor calls cudaTriggerProgrammaticLaunchCompletion() at its start, i.e. the dependent becomes
// eligible once the predecessor's last wave is resident); the dependent runs its prologue and then blocks in
// cudaGridDependencySynchronize() until the predecessor has completed and its writes are visible.  Hides the launch
// latency and prologue of k_assign and k_loss_stream (3 launches per step, ~180 us per step).
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

static int get_encoder(EncodeTiledFn &out) {
    static std::atomic<EncodeTiledFn> cached{nullptr};      // a process-wide driver entry point, not per-call state
    EncodeTiledFn encode = cached.load(std::memory_order_acquire);
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t ce = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        YS_REQUIRE(ce == cudaSuccess && fn != nullptr && q == cudaDriverEntryPointSuccess, YS_ERR_UNSUPPORTED,
                   "cuTensorMapEncodeTiled is not available from this driver");
        encode = (EncodeTiledFn)fn;
        cached.store(encode, std::memory_order_release);
    }
    out = encode;
    return YS_OK;
}
static int build_scan_maps(ScanMaps &M, const LevelSet &L, int B) {
    EncodeTiledFn encode;
    int rc = get_encoder(encode);
    if (rc) return rc;
    memset(&M, 0, sizeof(M));
    for (int l = 0; l < L.n; ++l) {
        if (!L.tma[l]) continue;
        const cuuint64_t gdim[2] = {(cuuint64_t)L.hw[l], (cuuint64_t)B * (cuuint64_t)L.C};
        const cuuint64_t gstride[1] = {(cuuint64_t)L.hw[l] * 4u};
        const cuuint32_t estr[2] = {1u, 1u};
        const cuuint32_t box[2] = {32u, 16u};       // one side (16 DFL bins) of one warp's 32 anchors
        CUresult r = encode(&M.lvl[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)L.x[l], gdim, gstride, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        YS_REQUIRE(r == CUDA_SUCCESS, YS_ERR_BAD_ARG, "tensor map (scan) of level %d: CUresult %d", l, (int)r);
    }
    return YS_OK;
}
static int build_stream_maps(StreamMaps &M, const LevelSet &L, int B) {
    EncodeTiledFn encode;
    int rc0 = get_encoder(encode);
    if (rc0) return rc0;
    memset(&M, 0, sizeof(M));
    for (int l = 0; l < L.n; ++l) {
        if (!L.tma[l]) continue;
        const cuuint64_t gdim[2] = {(cuuint64_t)L.hw[l], (cuuint64_t)B * (cuuint64_t)L.C};
        const cuuint64_t gstride[1] = {(cuuint64_t)L.hw[l] * 4u};
        const cuuint32_t estr[2] = {1u, 1u};
        const cuuint32_t box_b[2] = {(cuuint32_t)kBoxItem, 64u}, box_c[2] = {(cuuint32_t)kChunk, (cuuint32_t)kClsRows};
        CUresult r = encode(&M.box[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)L.x[l], gdim, gstride, box_b, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        YS_REQUIRE(r == CUDA_SUCCESS, YS_ERR_BAD_ARG, "tensor map (box) of level %d: CUresult %d", l, (int)r);
        r = encode(&M.cls[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)L.x[l], gdim, gstride, box_c, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        YS_REQUIRE(r == CUDA_SUCCESS, YS_ERR_BAD_ARG, "tensor map (cls) of level %d: CUresult %d", l, (int)r);
    }
    return YS_OK;
}

struct LossWorkspace {
    int *gt_off, *rank, *cursor, *gt_label, *owned_cnt, *gt_flags, *forced, *tgi, *lab, *img_order;
    unsigned char *risk;
    float4 *gt_box, *pbox, *partials;
    unsigned long long *best_key;
    float *norm, *tss, *scan_partial;
    double *img_sum;
    unsigned *pack_ctr;      // k_pack_gt (multi-block): [0] blocks finished, [1] "redo in one block" flag, [2 + b] launch weight of image b
    unsigned *counters;      // [0] images finished (k_assign), [1] blocks finished (k_loss), [2 + b] blocks of image b finished
    size_t bytes;
};

static inline int scan_blocks(int A) { return (A + kScanThreads - 1) / kScanThreads + YS_MAX_LEVELS; }   // upper bound (per-level cut)
// A = anchors per image as the caller counts them; per-anchor arrays are sized for the padded index space
static void carve_loss(LossWorkspace &W, void *base, size_t cap, int B, int A_true, int n_gt, int nc) {
    Carver c(base, cap);
    const size_t N = (size_t)(n_gt > 0 ? n_gt : 1);
    const size_t A = (size_t)padded_anchor_bound(A_true);
    W.counters = c.take<unsigned>((size_t)B + 2);
    W.pack_ctr = c.take<unsigned>((size_t)B + 2);
    W.tss = c.take<float>(64);
    W.gt_off = c.take<int>((size_t)B + 1);
    W.cursor = c.take<int>((size_t)B + 1);
    W.img_order = c.take<int>((size_t)B);
    W.rank = c.take<int>(N);
    W.gt_label = c.take<int>(N);
    W.owned_cnt = c.take<int>(N);
    W.gt_flags = c.take<int>(N);
    W.risk = c.take<unsigned char>(N);
    W.forced = c.take<int>(8 * N);
    W.gt_box = c.take<float4>(N);
    W.best_key = c.take<unsigned long long>(N);
    W.img_sum = c.take<double>((size_t)B);
    W.scan_partial = c.take<float>((size_t)B * scan_blocks((int)A));
    W.tgi = c.take<int>((size_t)B * A);
    W.lab = c.take<int>((size_t)B * A);
    W.norm = c.take<float>((size_t)B * A);
    W.pbox = c.take<float4>((size_t)B * A);
    W.partials = c.take<float4>(4096 + 2);      // one per k_loss_stream CTA (the grid is capped at 4096)
    W.bytes = align_up(c.off, 256);
}

}  // namespace ys

using namespace ys;

extern "C" size_t ys_detection_loss_workspace_bytes(int B, int A, int n_gt, int nc) {
    LossWorkspace W;
    carve_loss(W, nullptr, 0, B > 0 ? B : 1, A > 0 ? A : 1, n_gt, nc > 0 ? nc : 1);
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
    carve_loss(W, workspace, workspace_bytes, B, L.A_true, n_gt, nc);
    YS_REQUIRE(W.bytes <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu", W.bytes, workspace_bytes);
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    if (pred_boxes_px) { YS_REQUIRE(((uintptr_t)pred_boxes_px & 15) == 0, YS_ERR_BAD_ARG, "pred_boxes_px must be 16-byte aligned"); }
    if (target_gt_idx) { YS_REQUIRE(((uintptr_t)target_gt_idx & 15) == 0, YS_ERR_BAD_ARG, "target_gt_idx must be 16-byte aligned"); }
    if (norm_out) { YS_REQUIRE(((uintptr_t)norm_out & 15) == 0, YS_ERR_BAD_ARG, "norm must be 16-byte aligned"); }

    if ((stages & YS_STAGE_ALL) == 0) stages |= YS_STAGE_ALL;
    // k_assign's first wave of blocks (YS_SCAN_MINBLOCKS per SM) decodes before it may read k_pack_gt's image order; the
    // images of those launch slots are therefore fixed to 0..n_identity-1 (a function of the shapes only)
#ifndef YS_IDENTITY_WAVE
#define YS_IDENTITY_WAVE YS_SCAN_MINBLOCKS      // resident k_assign blocks per SM
#endif
    int n_identity = 0;
    int sms = 0;
    rc = device_sm_count(sms);
    if (rc) return rc;
    {
        const int wave = YS_IDENTITY_WAVE * sms;
        n_identity = (wave + L.n_sblk - 1) / L.n_sblk;
        if (n_identity > B) n_identity = B;
#ifdef YS_NO_EARLY       // experiment: every block waits for k_pack_gt first
        n_identity = 0;
#endif
    }
    if (stages & YS_STAGE_PACK) {
        if (n_gt <= kSmallMax && B <= kSmallMax)
            k_pack_gt_small<<<1, kSmallMax, 0, stream>>>(L, (const long long *)batch_idx, (const long long *)cls,
                                                        (const float4 *)bboxes, n_gt, B, nc, scale_x, scale_y, W.gt_off,
                                                        W.gt_box, W.gt_label, W.risk, W.best_key, W.owned_cnt,
                                                        W.gt_flags, W.counters, B + 2, W.img_order, n_identity);
        else {
            cudaError_t me = cudaMemsetAsync(W.pack_ctr, 0, ((size_t)B + 2) * sizeof(unsigned), stream);
            YS_REQUIRE(me == cudaSuccess, (int)me, "cudaMemsetAsync(pack_ctr): %s", cudaGetErrorString(me));
            const int blocks = n_gt > 0 ? (n_gt + kPackThreads - 1) / kPackThreads : 1;
            k_pack_gt<<<blocks, kPackThreads, 0, stream>>>(L, (const long long *)batch_idx, (const long long *)cls,
                                                           (const float4 *)bboxes, n_gt, B, nc, scale_x, scale_y, W.gt_off,
                                                           W.rank, W.cursor, W.gt_box, W.gt_label, W.risk, W.best_key,
                                                           W.owned_cnt, W.gt_flags, W.counters, B + 2, W.img_order,
                                                           n_identity, W.pack_ctr);
        }
        YS_CHECK_LAUNCH("k_pack_gt");
    }

    AssignIO io;
    io.gt_off = W.gt_off; io.G = 0; io.gt_box = W.gt_box; io.gt_label_i = W.gt_label; io.gt_label_f = nullptr;
    io.gt_valid = nullptr; io.pd_box = nullptr; io.pd_scores = nullptr; io.anc = nullptr;
    io.best_key = W.best_key; io.owned_cnt = W.owned_cnt; io.gt_flags = W.gt_flags; io.forced = W.forced;
    io.img_order = W.img_order;
#ifdef YS_NO_ORDER
    io.img_order = nullptr;
#endif
    // decode-before-wait is only sound behind this call's own k_pack_gt (an ordinary launch: everything earlier in the
    // stream, the producer of the head outputs included, has completed before it starts)
    io.n_identity = (stages & YS_STAGE_PACK) ? (io.img_order ? n_identity : B) : 0;
    io.skip_fixup = (stages & YS_STAGE_DIAG_NOFIXUP) ? 1 : 0;
    io.ciou = 0; io.sel = nullptr; io.sel_words = 0;
    // optional exports: written in place when the padded anchor space is the reference's (every h*w % 4 == 0), else
    // kept in the workspace and remapped by k_export_state after the assignment
    const bool remap = L.A != L.A_true;
    io.pbox = (pred_boxes_px && !remap) ? (float4 *)pred_boxes_px : W.pbox;
    io.tgi = (target_gt_idx && !remap) ? (int *)target_gt_idx : W.tgi;
    io.lab = W.lab;
    io.norm = (norm_out && !remap) ? norm_out : W.norm;
    io.partial = W.scan_partial; io.img_sum = W.img_sum; io.tss = W.tss;
    io.done_counter = W.counters; io.img_done = W.counters + 2;
    io.nc = nc; io.alpha = alpha; io.beta = beta; io.eps = 1e-9f;

    if (stages & YS_STAGE_ASSIGN) {
        ScanMaps SM;
        rc = build_scan_maps(SM, L, B);
        if (rc) return rc;
        const size_t smem = (size_t)kScanTileBytes + sizeof(AssignSmem);
        static DeviceOnce optin_assign;
        rc = optin_smem((const void *)k_assign<true>, smem, optin_assign, "k_assign");
        if (rc) return rc;
        dim3 gs(L.n_sblk, B);
        cudaError_t le = launch_pdl(k_assign<true>, gs, dim3(kScanThreads), smem, stream, L, io, B, SM);
        YS_REQUIRE(le == cudaSuccess, (int)le, "k_assign launch: %s", cudaGetErrorString(le));
        YS_CHECK_LAUNCH("k_assign");
        if (remap && (pred_boxes_px || target_gt_idx || norm_out)) {
            const size_t tot = (size_t)B * L.A_true;
            k_export_state<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(L, B, W.pbox, W.tgi, W.norm, (float4 *)pred_boxes_px,
                                                                             (int *)target_gt_idx, norm_out);
            YS_CHECK_LAUNCH("k_export_state");
        }
    }

    if (stages & YS_STAGE_LOSS) {
        // persistent TMA-fed stream: one CTA per SM
        StreamParams P;
        P.tgi = io.tgi; P.lab = W.lab; P.norm = io.norm; P.gt_off = W.gt_off; P.gt_box = W.gt_box; P.pbox = io.pbox;
        P.tss = W.tss; P.partials = W.partials; P.done = W.counters + 1; P.loss_out = loss_out;
        P.nc = nc; P.n_groups = (nc + kClsRows - 1) / kClsRows;
        P.n_box = L.n_hsegs;
        P.items_per_img = L.n_hsegs + L.n_chunks * P.n_groups;
        YS_REQUIRE((long long)B * P.items_per_img < 2147483647LL / 2, YS_ERR_UNSUPPORTED, "too many work items");
        P.total_items = B * P.items_per_img;
        P.ipi_magic = (unsigned)(0x100000000ull / (unsigned long long)P.items_per_img);
        P.nch_magic = (unsigned)(0x100000000ull / (unsigned long long)(L.n_chunks > 0 ? L.n_chunks : 1));
        if (P.items_per_img == 1) P.ipi_magic = 0xFFFFFFFFu;      // 2^32 does not fit: x * (2^32 - 1) >> 32 = x - 1, corrected once
        if (L.n_chunks == 1) P.nch_magic = 0xFFFFFFFFu;
        P.hyp_box = hyp_box; P.hyp_cls = hyp_cls; P.hyp_dfl = hyp_dfl; P.batch = (float)B;
        const int want = (P.total_items + kStreamWarps - 1) / kStreamWarps;
        const unsigned blocks = (unsigned)(want < sms ? (want > 0 ? want : 1) : (sms > 4096 ? 4096 : sms));
        // Ring depth: two stages where most box items are all-background (their neighbours in time are cheap and the SMs'
        // concurrent window of rows stays narrow: 3 % faster at the bench shape), three in dense scenes (tens of GT per
        // image and up: most items carry real work and a deeper ring keeps their tiles ahead: 11 % faster at C5)
#ifndef YS_STREAM_DENSE_GT
#define YS_STREAM_DENSE_GT 48
#endif
        const bool deep = (long long)n_gt >= (long long)YS_STREAM_DENSE_GT * B;
        const size_t smem = (size_t)kStreamWarps * (deep ? 3 : 2) * kStageBytes;
        {   // schedule (see k_loss_stream): rounds over contiguous windows of `blocks` groups, rotated per round so that a
            // CTA's position inside an image drifts by ~0.382 of the image's items per round
            const long long NG = ((long long)P.total_items + kStreamWarps - 1) / kStreamWarps;
            const long long G = (long long)blocks, ipi = (long long)P.items_per_img;
            P.n_groups16 = (unsigned)NG;
            P.rounds = (unsigned)((NG + G - 1) / G);
            const long long d0 = (G * kStreamWarps) % ipi;                 // natural drift per round, in items
#ifndef YS_STREAM_DRIFT
#define YS_STREAM_DRIFT 0.5
#endif
            long long want_drift = (long long)(YS_STREAM_DRIFT * (double)ipi) - d0;
            want_drift %= ipi;
            if (want_drift < 0) want_drift += ipi;
            long long r = (want_drift + kStreamWarps / 2) / kStreamWarps;   // rotation in groups
#ifdef YS_STREAM_ROTATE
            r = YS_STREAM_ROTATE;
#endif
            P.rotate = (unsigned)(G > 1 ? r % G : 0);
        }
        StreamMaps M;
        rc = build_stream_maps(M, L, B);
        if (rc) return rc;
        static DeviceOnce optin_ls[4];
        typedef void (*StreamKernel)(LevelSet, StreamParams, const StreamMaps);
        const int kid = (level_grad_host ? 0 : 2) + (deep ? 1 : 0);
        const StreamKernel kern = level_grad_host ? (deep ? (StreamKernel)k_loss_stream<true, 3> : (StreamKernel)k_loss_stream<true, 2>)
                                                  : (deep ? (StreamKernel)k_loss_stream<false, 3> : (StreamKernel)k_loss_stream<false, 2>);
        rc = optin_smem((const void *)kern, smem, optin_ls[kid], "k_loss_stream");
        if (rc) return rc;
        cudaError_t le = launch_pdl(kern, dim3(blocks), dim3(kStreamThreads), smem, stream, L, P, M);
        YS_REQUIRE(le == cudaSuccess, (int)le, "k_loss_stream launch: %s", cudaGetErrorString(le));
        YS_CHECK_LAUNCH("k_loss_stream");
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
    float *norm, *tss, *scan_partial;
    double *img_sum;
    unsigned *sel;       // top-k mode: selection bitmask (B*G rows of ceil(A/32) words)
    unsigned *counters;  // [0] images finished, [2 + b] blocks of image b finished
    size_t zero_bytes;   // leading region that must be zeroed per call
    size_t bytes;
};
static void carve_assign(AssignWorkspace &W, void *base, size_t cap, int B, int A, int G, int topk = 0) {
    Carver c(base, cap);
    const size_t N = (size_t)B * G;
    W.counters = c.take<unsigned>((size_t)B + 2);
    W.owned_cnt = c.take<int>(N);
    W.gt_flags = c.take<int>(N);
    W.zero_bytes = align_up(c.off, 256);
    W.best_key = c.take<unsigned long long>(N);
    W.tss = c.take<float>(64);
    W.forced = c.take<int>(8 * N);
    W.img_sum = c.take<double>((size_t)B);
    W.scan_partial = c.take<float>((size_t)B * scan_blocks(A));
    W.tgi = c.take<int>((size_t)B * A);
    W.norm = c.take<float>((size_t)B * A);
    W.sel = topk > 0 ? c.take<unsigned>(N * (size_t)((A + 31) / 32)) : nullptr;
    W.bytes = align_up(c.off, 256);
}
}  // namespace ys

extern "C" size_t ys_assign_workspace_bytes(int B, int A, int G) {
    AssignWorkspace W;
    carve_assign(W, nullptr, 0, B > 0 ? B : 1, A > 0 ? A : 1, G > 0 ? G : 1);
    return W.bytes;
}
extern "C" size_t ys_assign_ex_workspace_bytes(int B, int A, int G, int topk) {
    AssignWorkspace W;
    carve_assign(W, nullptr, 0, B > 0 ? B : 1, A > 0 ? A : 1, G > 0 ? G : 1, topk);
    return W.bytes;
}

extern "C" int ys_assign(int B, int A, int G, int nc, const float *pd_scores, const float *pd_bboxes,
                         const float *anc_points, const float *gt_labels, const float *gt_bboxes,
                         const float *mask_gt, float alpha, float beta, float eps, float *target_labels,
                         float *target_bboxes, float *target_scores, uint8_t *fg_mask, int64_t *target_gt_idx,
                         float *norm_out, void *workspace, size_t workspace_bytes, void *stream_v) {
    return ys_assign_ex(B, A, G, nc, pd_scores, pd_bboxes, anc_points, gt_labels, gt_bboxes, mask_gt, alpha, beta, eps, 0, 0,
                        target_labels, target_bboxes, target_scores, fg_mask, target_gt_idx, norm_out, workspace,
                        workspace_bytes, stream_v);
}

extern "C" int ys_assign_ex(int B, int A, int G, int nc, const float *pd_scores, const float *pd_bboxes,
                            const float *anc_points, const float *gt_labels, const float *gt_bboxes,
                            const float *mask_gt, float alpha, float beta, float eps, int topk, int ciou_assign,
                            float *target_labels, float *target_bboxes, float *target_scores, uint8_t *fg_mask,
                            int64_t *target_gt_idx, float *norm_out, void *workspace, size_t workspace_bytes,
                            void *stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    YS_REQUIRE(topk >= 0, YS_ERR_BAD_ARG, "topk must be >= 0 (0 = off, the reference's behaviour)");
    YS_REQUIRE(topk == 0 || G <= 65535, YS_ERR_UNSUPPORTED, "top-k mode: G must be <= 65535 (grid.x)");
    YS_REQUIRE(B >= 1 && A >= 1 && G >= 1 && nc >= 1, YS_ERR_BAD_ARG, "B, A, G, nc must be >= 1");
    YS_REQUIRE(B <= 65535, YS_ERR_UNSUPPORTED, "B must be <= 65535 (grid.y)");
    YS_REQUIRE(pd_scores && pd_bboxes && anc_points && gt_labels && gt_bboxes && mask_gt, YS_ERR_BAD_ARG, "null input");
    YS_REQUIRE(target_labels && target_bboxes && fg_mask && target_gt_idx && workspace, YS_ERR_BAD_ARG, "null output");
    YS_REQUIRE((((uintptr_t)pd_bboxes | (uintptr_t)gt_bboxes | (uintptr_t)target_bboxes) & 15) == 0 &&
                   ((uintptr_t)anc_points & 7) == 0, YS_ERR_BAD_ARG, "box pointers must be 16-byte aligned, anc_points 8-byte");
    AssignWorkspace W;
    carve_assign(W, workspace, workspace_bytes, B, A, G, topk);
    YS_REQUIRE(W.bytes <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu", W.bytes, workspace_bytes);
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    cudaMemsetAsync(workspace, 0, W.zero_bytes, stream);

    LevelSet L;
    L.n = 1; L.A = A; L.A_true = A; L.C = nc; L.n_chunks = 0;      // dense entry point: one index space, no levels
    for (int i = 0; i < YS_MAX_LEVELS; ++i) { L.x[i] = nullptr; L.g[i] = nullptr; L.w[i] = 1; L.hw[i] = 0; L.stride[i] = 1.f; L.tma[i] = 0; }
    for (int i = 0; i <= YS_MAX_LEVELS; ++i) { L.a_off[i] = i == 0 ? 0 : A; L.a_true_off[i] = L.a_off[i]; L.chunk_off[i] = 0; L.hseg_off[i] = 0; L.sblk_off[i] = 0; }
    L.n_hsegs = 0; L.n_sblk = 0;

    AssignIO io;
    io.gt_off = nullptr; io.G = G; io.gt_box = (const float4 *)gt_bboxes; io.gt_label_i = nullptr;
    io.gt_label_f = gt_labels; io.gt_valid = mask_gt; io.pd_box = (const float4 *)pd_bboxes; io.pd_scores = pd_scores;
    io.anc = (const float2 *)anc_points; io.best_key = W.best_key; io.owned_cnt = W.owned_cnt; io.gt_flags = W.gt_flags;
    io.img_order = nullptr; io.n_identity = 0; io.skip_fixup = 0;
    io.ciou = ciou_assign ? 1 : 0; io.sel = nullptr; io.sel_words = (A + 31) / 32;
    if (topk > 0) {
        k_topk_select<<<dim3(G, B), kTopkThreads, 0, stream>>>(A, G, nc, pd_scores, (const float4 *)pd_bboxes,
                                                               (const float2 *)anc_points, gt_labels,
                                                               (const float4 *)gt_bboxes, mask_gt, alpha, beta, io.ciou, topk,
                                                               W.sel, io.sel_words);
        YS_CHECK_LAUNCH("k_topk_select");
        io.sel = W.sel;
    }
    io.forced = W.forced; io.pbox = nullptr; io.tgi = W.tgi; io.lab = nullptr; io.norm = norm_out ? norm_out : W.norm;
    io.partial = W.scan_partial; io.img_sum = W.img_sum; io.tss = W.tss;
    io.done_counter = W.counters; io.img_done = W.counters + 2;
    io.nc = nc; io.alpha = alpha; io.beta = beta; io.eps = eps;

    ScanMaps SM;
    memset(&SM, 0, sizeof(SM));
    dim3 gs((A + kScanThreads - 1) / kScanThreads, B);
    k_assign<false><<<gs, kScanThreads, sizeof(AssignSmem), stream>>>(L, io, B, SM);
    YS_CHECK_LAUNCH("k_assign(dense)");
    const size_t tot = (size_t)B * A;
    k_assign_emit<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(B, A, G, nc, W.tgi, io.norm, gt_labels,
                                                                   (const float4 *)gt_bboxes, target_labels,
                                                                   (float4 *)target_bboxes, target_scores, fg_mask,
                                                                   (long long *)target_gt_idx);
    YS_CHECK_LAUNCH("k_assign_emit");
    return YS_OK;
}
