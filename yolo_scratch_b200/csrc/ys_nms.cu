// ys_nms.cu -- batched greedy NMS (bit-exact keep lists) for the validation path.
// Replaces decode_predictions_for_metrics / nms_simple / calculate_iou_batch_simple
// (train_yolo11_cuda.py:265-437 of the reference); see include/ys_b200.h.
//
//   k_nms_filter    streaming pass over pred: best class (first max), score > conf, xywh -> xyxy,
//                   sort key (score bits << 32 | ~anchor), warp-aggregated compaction per image
//   k_nms_sort      one block per image: bitonic sort of the keys in shared memory (descending; ties ->
//                   lower anchor index first), gathers boxes / classes into sorted order
//   k_nms_suppress  one block per image: sorted boxes resident in shared memory; 64-row tiles; the
//                   64x64 diagonal bitmask tile is built with warp ballots and resolved serially, then the
//                   kept rows of the tile suppress all later columns in parallel (no N x N matrix in HBM)
#include "ys_common.cuh"

namespace ys {

constexpr int kFilterThreads = 256;
constexpr int kSortThreads = 1024;
constexpr int kSupThreads = 512;
constexpr int kSortMaxN = 16384;       // keys per image the shared-memory sort handles (128 KB)
constexpr int kSupSmemBoxes = 11264;   // boxes per image kept resident in shared memory (176 KB + extras)

__device__ __forceinline__ unsigned score_bits_ordered(float s) {
    unsigned u = __float_as_uint(s);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);   // order-preserving for all finite floats; never 0 for s > -NaN
}
__device__ __forceinline__ float score_from_ordered(unsigned o) {
    unsigned u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
    return __uint_as_float(u);
}

// ---------------------------------------------------------------------------------------------------
// filter.  layout 1: pred (B, A, R) rows of R = 4+nc floats; layout 0: pred (B, R, A).
// One warp handles 32 consecutive anchors.
// ---------------------------------------------------------------------------------------------------
template <int LAYOUT>
__global__ void __launch_bounds__(kFilterThreads)
k_nms_filter(const float *__restrict__ pred, int B, int A, int nc, float conf, unsigned long long *__restrict__ keys,
             float4 *__restrict__ boxes, int *__restrict__ cls, int *__restrict__ count) {
    extern __shared__ __align__(16) float s_rows[];   // LAYOUT 1: per warp 32 rows with odd stride
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps_per_img = (A + 31) / 32;
    const long long gw = (long long)blockIdx.x * (kFilterThreads / 32) + warp;
    if (gw >= (long long)B * warps_per_img) return;
    const int b = (int)(gw / warps_per_img);
    const int a0 = (int)(gw % warps_per_img) * 32;
    const int a = a0 + lane;
    const int R = 4 + nc;
    float cx = 0.f, cy = 0.f, bw = 0.f, bh = 0.f, best = 0.f;
    int bc = 0;
    if (LAYOUT == 1 && (R & 3) == 0 && (reinterpret_cast<uintptr_t>(pred) & 15) == 0) {
        // Rows of R floats, R a multiple of 4: the warp's 32 rows are one contiguous, 16-byte aligned run of 32*R*4
        // bytes.  It is copied as float4 (all of a lane's loads in flight together) into shared memory at the same
        // linear offsets, and every lane then walks ITS row as float4: row stride R/4 float4 (21 at nc = 80, odd ->
        // the eight lanes of a quarter-warp hit eight different 16-byte bank groups).  The scalar version below
        // (one division per element, 4-byte loads) ran at 0.78 TB/s.
        const int R4 = R >> 2;
        float4 *mine4 = reinterpret_cast<float4 *>(s_rows + (size_t)warp * 32 * (R | 1));      // 128 * (R|1) bytes per warp
        const int rows = min(32, A - a0);
        const float4 *src4 = reinterpret_cast<const float4 *>(pred + ((size_t)b * A + a0) * R);
        const int total4 = rows * R4;
        // cp.async: all of the warp's 16-byte pieces in flight at once, no register holds a load (the __ldg + st.shared
        // version kept 7 per lane in flight: 4.1 TB/s)
        for (int i = lane; i < total4; i += 32) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(mine4 + i);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src4 + i) : "memory");
        }
        asm volatile("cp.async.commit_group;\n cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        if (a < A) {
            const float4 *row4 = mine4 + lane * R4;
            const float4 q = row4[0];
            cx = q.x; cy = q.y; bw = q.z; bh = q.w;
            best = row4[1].x;                                           // row[4]: the starting value (class 0)
            for (int j = 1; j < R4; ++j) {
                const float4 v = row4[j];
                const int c = (j - 1) * 4;
                if (v.x > best) { best = v.x; bc = c; }
                if (v.y > best) { best = v.y; bc = c + 1; }             // first maximum (torch.max(dim) semantics)
                if (v.z > best) { best = v.z; bc = c + 2; }
                if (v.w > best) { best = v.w; bc = c + 3; }
            }
        }
    } else if (LAYOUT == 1) {
        const int RS = R | 1;
        float *mine = s_rows + (size_t)warp * 32 * RS;
        const int rows = min(32, A - a0);
        const float *src = pred + ((size_t)b * A + a0) * R;
        const int total = rows * R;
        for (int i = lane; i < total; i += 32) {
            int r = i / R, c = i - r * R;
            mine[r * RS + c] = __ldg(src + i);
        }
        __syncwarp();
        if (a < A) {
            const float *row = mine + lane * RS;
            cx = row[0]; cy = row[1]; bw = row[2]; bh = row[3];
            best = row[4];
            for (int c = 1; c < nc; ++c) {
                float v = row[4 + c];
                if (v > best) { best = v; bc = c; }     // first maximum (torch.max(dim) semantics)
            }
        }
    } else {
        if (a < A) {
            const float *col = pred + (size_t)b * R * A + a;
            cx = __ldg(col); cy = __ldg(col + A); bw = __ldg(col + 2 * (size_t)A); bh = __ldg(col + 3 * (size_t)A);
            best = __ldg(col + 4 * (size_t)A);
            for (int c = 1; c < nc; ++c) {
                float v = __ldg(col + (size_t)(4 + c) * A);
                if (v > best) { best = v; bc = c; }
            }
        }
    }
    const bool pass = (a < A) && (best > conf);                         // train_yolo11_cuda.py:307
    const unsigned m = __ballot_sync(0xffffffffu, pass);
    if (m == 0) return;
    int base = 0;
    if (lane == 0) base = atomicAdd(&count[b], __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (pass) {
        const int pos = base + __popc(m & ((1u << lane) - 1u));
        // x - w/2 ... (train_yolo11_cuda.py:324-329); /2 is exact
        const float hw = bw * 0.5f, hh = bh * 0.5f;
        keys[(size_t)b * A + pos] = ((unsigned long long)score_bits_ordered(best) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)a);
        boxes[(size_t)b * A + a] = make_float4(YS_SUB(cx, hw), YS_SUB(cy, hh), YS_ADD(cx, hw), YS_ADD(cy, hh));
        cls[(size_t)b * A + a] = bc;
    }
}

// keys for nms_single: every input box is a candidate
__global__ void k_nms_keys_single(const float *__restrict__ scores, int N, unsigned long long *__restrict__ keys, int *__restrict__ count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) keys[i] = ((unsigned long long)score_bits_ordered(scores[i]) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)i);
    if (i == 0) count[0] = N;
}

// ---------------------------------------------------------------------------------------------------
// sort: descending order of count[b] keys (unique: score bits << 32 | ~anchor) in shared memory.
// A bitonic network needs a power-of-two length, and 8400 candidates would mean 16384 slots, half of them padding
// (105 stages of 8 compare-exchanges per thread).  Instead the keys are split at H = the largest power of two <= n:
// the first H are sorted by one network, the remaining n - H (padded to their own power of two) by a second, small
// one, and every key then finds its final position by one binary search in the OTHER run (own index + number of
// larger keys there) and is gathered straight to its place in the sorted outputs: 8192 + 256 slots instead of 16384
// at n = 8400.
// ---------------------------------------------------------------------------------------------------
// Plain network, one compare-exchange per thread and step, a block barrier per step: runs shorter than 128 keys only.
__device__ __forceinline__ void nms_bitonic_small(unsigned long long *s, int P) {      // whole block; ends on a barrier
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (P >> 1); t += kSortThreads) {
                // t-th compare-exchange of this stage
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int p = i | j;
                const bool desc = ((i & k) == 0);
                const unsigned long long x = s[i], y = s[p];
                if ((x < y) == desc) { s[i] = y; s[p] = x; }
            }
            __syncthreads();
        }
    }
}

// The same network for P >= 128 with the keys in REGISTERS between steps (the plain version moves every key through
// shared memory once per step -- 91 steps for 8192 keys: 221 us for 256 images; this one 160 us, bound by the 64-bit
// compare / select instructions).  Two kinds of passes, each one load and one store of the run:
//   chunk pass  a warp owns 128 consecutive slots, lane holds slots lane + 32 m (m = 0..3): the steps with distance 64
//               and 32 are compare-exchanges inside the thread, 16..1 go by shuffle.  The first pass runs all 28 steps of
//               the sizes k = 2..128 this way; afterwards every size k ends with one (steps 64..1).
//   cross pass  up to three steps with distances j, j/2, j/4 >= 128: a thread holds the 8 slots base + a * (j/4), so all
//               three are compare-exchanges inside the thread; consecutive threads hold consecutive slots.
// 16 passes instead of 91 at 8192 keys.  Same comparisons as the plain network, hence the same (unique) result.
__device__ __forceinline__ void nms_ce(unsigned long long &a, unsigned long long &b, bool desc) {
    if ((a < b) == desc) { const unsigned long long t = a; a = b; b = t; }
}
// steps j_hi, j_hi/2, .., 1 (j_hi <= 64) of size k on the four keys of this lane; idx0 = slot of x[0]
__device__ __forceinline__ void nms_chunk_steps(unsigned long long (&x)[4], int idx0, int lane, int k, int j_hi) {
    if (j_hi >= 64) {
        nms_ce(x[0], x[2], (idx0 & k) == 0);
        nms_ce(x[1], x[3], ((idx0 + 32) & k) == 0);
    }
    if (j_hi >= 32) {
        nms_ce(x[0], x[1], (idx0 & k) == 0);
        nms_ce(x[2], x[3], ((idx0 + 64) & k) == 0);
    }
    for (int j = j_hi < 16 ? j_hi : 16; j > 0; j >>= 1) {
        const bool lower = (lane & j) == 0;
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            const unsigned long long y = __shfl_xor_sync(0xffffffffu, x[m], j);
            const bool desc = ((idx0 + 32 * m) & k) == 0;        // bit k of a slot and of its partner agree (k > j)
            const bool want_max = lower == desc;                 // descending: the lower slot keeps the larger key
            x[m] = ((x[m] > y) == want_max) ? x[m] : y;          // keys are unique (equal padding zeros: either one)
        }
    }
}
__device__ __forceinline__ void nms_bitonic_desc(unsigned long long *s, int P) {      // whole block; ends on a barrier
    if (P < 128) { nms_bitonic_small(s, P); return; }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // ---- sizes 2..128: everything inside the 128-slot chunks
    for (int c = warp; c < (P >> 7); c += kSortThreads / 32) {
        const int idx0 = (c << 7) + lane;
        unsigned long long x[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) x[m] = s[idx0 + 32 * m];
        for (int k = 2; k <= 128; k <<= 1) nms_chunk_steps(x, idx0, lane, k, k >> 1);
#pragma unroll
        for (int m = 0; m < 4; ++m) s[idx0 + 32 * m] = x[m];
    }
    __syncthreads();
    for (int k = 256; k <= P; k <<= 1) {
        // ---- cross passes: distances k/2 .. 128, three steps at a time
        int j = k >> 1;
        while (j >= 128) {
            int S = 1;
            if (j >= 256) S = 2;
            if (j >= 512) S = 3;
            const int jl = j >> (S - 1);                           // smallest distance of this pass
            for (int t = threadIdx.x; t < (P >> S); t += kSortThreads) {
                const int base = ((t & ~(jl - 1)) << S) | (t & (jl - 1));
                const bool desc = (base & k) == 0;                 // k > j: the same for all slots of the thread
                unsigned long long x[8];
#pragma unroll
                for (int a = 0; a < 8; ++a)
                    if (a < (1 << S)) x[a] = s[base + a * jl];
#pragma unroll
                for (int st = 2; st >= 0; --st) {
                    if (st < S) {
#pragma unroll
                        for (int a = 0; a < 8; ++a)
                            if (a < (1 << S) && !(a & (1 << st))) nms_ce(x[a], x[a | (1 << st)], desc);
                    }
                }
#pragma unroll
                for (int a = 0; a < 8; ++a)
                    if (a < (1 << S)) s[base + a * jl] = x[a];
            }
            __syncthreads();
            j >>= S;
        }
        // ---- distances 64..1 inside the chunks
        for (int c = warp; c < (P >> 7); c += kSortThreads / 32) {
            const int idx0 = (c << 7) + lane;
            unsigned long long x[4];
#pragma unroll
            for (int m = 0; m < 4; ++m) x[m] = s[idx0 + 32 * m];
            nms_chunk_steps(x, idx0, lane, k, 64);
#pragma unroll
            for (int m = 0; m < 4; ++m) s[idx0 + 32 * m] = x[m];
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kSortThreads)      // 50 registers; capped at 32 for two blocks per SM it spills and is no faster
k_nms_sort(const unsigned long long *__restrict__ keys, const int *__restrict__ count, int A,
           const float4 *__restrict__ boxes, const int *__restrict__ cls, float4 *__restrict__ sbox,
           float *__restrict__ sscore, int *__restrict__ scls, int *__restrict__ sanchor) {
    extern __shared__ unsigned long long s_keys[];
    const int b = blockIdx.x;
    const int n = count[b];
    if (n == 0) return;
    int H = 1;
    while (H * 2 <= n) H <<= 1;                  // largest power of two <= n
    const int nb = n - H;                        // keys of the second run
    int P2 = 0;
    if (nb > 0) { P2 = 1; while (P2 < nb) P2 <<= 1; }
    unsigned long long *run_a = s_keys, *run_b = s_keys + H;      // H + P2 <= the power of two >= n the host sized for
    const unsigned long long *src = keys + (size_t)b * A;
    for (int i = threadIdx.x; i < H + P2; i += kSortThreads) s_keys[i] = i < n ? src[i] : 0ull;      // real keys are never 0
    __syncthreads();
    nms_bitonic_desc(run_a, H);
    if (P2 > 1) nms_bitonic_desc(run_b, P2);
    for (int r = threadIdx.x; r < n; r += kSortThreads) {
        const unsigned long long key = s_keys[r];
        // how many keys of the other run are larger (both runs descend; keys are unique)
        const unsigned long long *other = r < H ? run_b : run_a;
        int lo = 0, hi = r < H ? nb : H;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (other[mid] > key) lo = mid + 1; else hi = mid;
        }
        const int pos = (r < H ? r : r - H) + lo;
        const int a = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
        const size_t o = (size_t)b * A + pos;
        sbox[o] = boxes[(size_t)b * A + a];
        sscore[o] = score_from_ordered((unsigned)(key >> 32));
        scls[o] = cls ? cls[(size_t)b * A + a] : 0;
        sanchor[o] = a;
    }
}

// ---------------------------------------------------------------------------------------------------
// sort of more than kSortMaxN keys per image (1280 px validation: 33600 candidates; the reference's nms_simple has no
// limit, train_yolo11_cuda.py:361-399).  The image's keys are cut into segments of kSortMaxN: k_nms_sort_seg sorts every
// segment in shared memory (one block per segment, descending) and writes it back; k_nms_merge then gives every key its
// final rank = its index inside its own segment + the number of larger keys in every other segment (one binary search
// per other segment; keys are unique) and gathers the box / class / anchor into sorted order, as k_nms_sort does.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
k_nms_sort_seg(const unsigned long long *__restrict__ keys, const int *__restrict__ count, int A,
               unsigned long long *__restrict__ keys_sorted) {
    extern __shared__ unsigned long long s_keys[];
    const int b = blockIdx.y, seg = blockIdx.x;
    const int n = count[b];
    const int lo = seg * kSortMaxN;
    if (lo >= n) return;
    const int m = min(kSortMaxN, n - lo);
    int P = 1;
    while (P < m) P <<= 1;
    const unsigned long long *src = keys + (size_t)b * A + lo;
    for (int i = threadIdx.x; i < P; i += kSortThreads) s_keys[i] = i < m ? src[i] : 0ull;      // real keys are never 0
    __syncthreads();
    if (P > 1) nms_bitonic_desc(s_keys, P);
    unsigned long long *dst = keys_sorted + (size_t)b * A + lo;
    for (int i = threadIdx.x; i < m; i += kSortThreads) dst[i] = s_keys[i];
}

__global__ void __launch_bounds__(256)
k_nms_merge(const unsigned long long *__restrict__ keys_sorted, const int *__restrict__ count, int A,
            const float4 *__restrict__ boxes, const int *__restrict__ cls, float4 *__restrict__ sbox,
            float *__restrict__ sscore, int *__restrict__ scls, int *__restrict__ sanchor) {
    const int b = blockIdx.y;
    const int n = count[b];
    const int r = blockIdx.x * 256 + threadIdx.x;
    if (r >= n) return;
    const unsigned long long *base = keys_sorted + (size_t)b * A;
    const unsigned long long key = base[r];
    const int seg = r / kSortMaxN;
    int pos = r - seg * kSortMaxN;
    const int nseg = (n + kSortMaxN - 1) / kSortMaxN;
    for (int s2 = 0; s2 < nseg; ++s2) {
        if (s2 == seg) continue;
        const unsigned long long *other = base + (size_t)s2 * kSortMaxN;
        int lo = 0, hi = min(kSortMaxN, n - s2 * kSortMaxN);
        while (lo < hi) {                         // number of keys of segment s2 larger than this one (it descends)
            const int mid = (lo + hi) >> 1;
            if (other[mid] > key) lo = mid + 1; else hi = mid;
        }
        pos += lo;
    }
    const int a = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
    const size_t o = (size_t)b * A + pos;
    sbox[o] = boxes[(size_t)b * A + a];
    sscore[o] = score_from_ordered((unsigned)(key >> 32));
    scls[o] = cls ? cls[(size_t)b * A + a] : 0;
    sanchor[o] = a;
}

// ---------------------------------------------------------------------------------------------------
// suppress
// ---------------------------------------------------------------------------------------------------
// IoU(i, j) > thr with the reference's arithmetic (train_yolo11_cuda.py:418-435, :396).
__device__ __forceinline__ bool suppresses(const float4 &bi, float area_i, const float4 &bj, float thr) {
    const float iw = YS_SUB(fminf2(bi.z, bj.z), fmaxf2(bi.x, bj.x));
    const float ih = YS_SUB(fminf2(bi.w, bj.w), fmaxf2(bi.y, bj.y));
    if (!(iw > 0.f) || !(ih > 0.f)) {
        // inter == 0 (or NaN): IoU = 0/(union+1e-6).  Only a non-positive denominator could make that
        // exceed thr; fall through to the exact path in that (degenerate-box) case.
        if (iw == iw && ih == ih) {
            const float area_j = area_plain(bj);
            const float uni0 = YS_ADD(YS_ADD(area_i, area_j), kEpsNms);   // inter = 0: (a1+a2-0)+eps
            if (uni0 > 0.f) return 0.f > thr;
        }
    }
    const float area_j = area_plain(bj);
    const float iou = iou_nms(bi, area_i, bj, area_j);
    return !(iou <= thr);   // the reference keeps `ious <= thr`; NaN is dropped (:396)
}

template <bool CLASS_AWARE>
__global__ void __launch_bounds__(kSupThreads)
k_nms_suppress(const int *__restrict__ count, int A, const float4 *__restrict__ sbox, const float *__restrict__ sscore,
               const int *__restrict__ scls, const int *__restrict__ sanchor, float thr, float img_size, int smem_boxes,
               int *__restrict__ keep_anchor, float4 *__restrict__ keep_boxes, float *__restrict__ keep_scores,
               int *__restrict__ keep_labels, int *__restrict__ keep_count, const int *__restrict__ only) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int b = blockIdx.x;
    if (only && !only[b]) return;      // k_nms_suppress_grid has done this image
    const int n = count[b];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (n == 0) {
        if (tid == 0) keep_count[b] = 0;
        return;
    }
    const int nwords = (n + 31) / 32;
    // shared layout: boxes[smem_boxes] (float4) | removed[nwords32] | diag[64] (u64) | tile bookkeeping
    float4 *s_box = reinterpret_cast<float4 *>(s_raw);
    const bool resident = n <= smem_boxes;
    const int nb = resident ? n : 0;
    unsigned *s_removed = reinterpret_cast<unsigned *>(s_raw + (size_t)(resident ? smem_boxes : 0) * sizeof(float4));
    const int max_words = (A + 31) / 32;
    unsigned long long *s_diag = reinterpret_cast<unsigned long long *>(s_removed + ((max_words + 3) & ~3));
    float4 *s_tile = reinterpret_cast<float4 *>(s_diag + 64);   // 64 boxes of the current row tile
    float *s_tarea = reinterpret_cast<float *>(s_tile + 64);
    int *s_tcls = reinterpret_cast<int *>(s_tarea + 64);
    int *s_kept = s_tcls + 64;                                  // [0] = number kept in tile, [1..64] their tile rows
    const float4 *gbox = sbox + (size_t)b * A;
    const int *gcls = scls + (size_t)b * A;

    for (int i = tid; i < nb; i += kSupThreads) s_box[i] = gbox[i];
    for (int i = tid; i < nwords; i += kSupThreads) s_removed[i] = 0u;
    __syncthreads();
    const float4 *boxp = resident ? s_box : gbox;

    for (int t0 = 0; t0 < n; t0 += 64) {
        const int tn = min(64, n - t0);
        if (tid < 64) {
            float4 bx = tid < tn ? boxp[t0 + tid] : make_float4(0.f, 0.f, 0.f, 0.f);
            s_tile[tid] = bx;
            s_tarea[tid] = area_plain(bx);
            if (CLASS_AWARE) s_tcls[tid] = tid < tn ? gcls[t0 + tid] : -1;
        }
        __syncthreads();
        // -- diagonal 64x64 bitmask tile by warp ballots: warp w handles rows w, w+16, w+32, w+48
        for (int r = warp; r < tn; r += kSupThreads / 32) {
            const float4 bi = s_tile[r];
            const float ai = s_tarea[r];
            unsigned long long bits = 0ull;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = h * 32 + lane;
                bool sup = false;
                if (c > r && c < tn) {
                    sup = suppresses(bi, ai, s_tile[c], thr);
                    if (CLASS_AWARE) sup = sup && (s_tcls[c] == s_tcls[r]);
                }
                const unsigned m = __ballot_sync(0xffffffffu, sup);
                bits |= (unsigned long long)m << (32 * h);
            }
            if (lane == 0) s_diag[r] = bits;
        }
        __syncthreads();
        // -- serial resolution of the tile (one thread): kept rows and the rows they remove inside the tile
        if (tid == 0) {
            const int w0 = t0 >> 5;
            unsigned long long rem = (unsigned long long)s_removed[w0];
            if (tn > 32) rem |= (unsigned long long)s_removed[w0 + 1] << 32;
            int nk = 0;
            for (int r = 0; r < tn; ++r) {
                if (!((rem >> r) & 1ull)) {
                    s_kept[1 + nk++] = r;
                    rem |= s_diag[r];
                }
            }
            s_kept[0] = nk;
            s_removed[w0] = (unsigned)(rem & 0xFFFFFFFFull);
            if (tn > 32) s_removed[w0 + 1] = (unsigned)(rem >> 32);
        }
        __syncthreads();
        // -- kept rows of this tile suppress every later, still-alive column
        const int nk = s_kept[0];
        if (nk > 0) {
            for (int c = t0 + 64 + tid; c < n; c += kSupThreads) {
                if ((s_removed[c >> 5] >> (c & 31)) & 1u) continue;
                const float4 bj = boxp[c];
                const int cj = CLASS_AWARE ? gcls[c] : 0;
                bool dead = false;
                for (int q = 0; q < nk && !dead; ++q) {
                    const int r = s_kept[1 + q];
                    if (CLASS_AWARE && s_tcls[r] != cj) continue;
                    dead = suppresses(s_tile[r], s_tarea[r], bj, thr);
                }
                if (dead) atomicOr(&s_removed[c >> 5], 1u << (c & 31));
            }
        }
        __syncthreads();
    }

    // -- compact the survivors in sorted order (one warp, ballot prefix)
    if (warp == 0) {
        int base = 0;
        for (int c0 = 0; c0 < n; c0 += 32) {
            const int c = c0 + lane;
            const bool alive = c < n && !((s_removed[c >> 5] >> (c & 31)) & 1u);
            const unsigned m = __ballot_sync(0xffffffffu, alive);
            if (alive) {
                const size_t o = (size_t)b * A + base + __popc(m & ((1u << lane) - 1u));
                const size_t s = (size_t)b * A + c;
                if (keep_anchor) keep_anchor[o] = sanchor[s];
                if (keep_boxes) {
                    const float4 bx = sbox[s];
                    // boxes / img_size, clamp to [0,1]  (train_yolo11_cuda.py:347-350)
                    float4 q = make_float4(YS_DIV(bx.x, img_size), YS_DIV(bx.y, img_size), YS_DIV(bx.z, img_size), YS_DIV(bx.w, img_size));
                    q.x = fminf2(fmaxf2(q.x, 0.f), 1.f); q.y = fminf2(fmaxf2(q.y, 0.f), 1.f);
                    q.z = fminf2(fmaxf2(q.z, 0.f), 1.f); q.w = fminf2(fmaxf2(q.w, 0.f), 1.f);
                    keep_boxes[o] = q;
                }
                if (keep_scores) keep_scores[o] = sscore[s];
                if (keep_labels) keep_labels[o] = scls[s];
            }
            base += __popc(m);
        }
        if (lane == 0) keep_count[b] = base;
    }
}

constexpr int kRegThreads = 1024;

// Half of the overlap a pair needs along one axis before IoU > thr is possible, charged to one box of extent [lo, hi]:
// IoU > thr  =>  inter > thr * union >= thr * area_j  =>  iw * ih > thr * w_j * h_j, and ih <= h_j  =>  iw > thr * w_j
// (likewise for box i), hence iw > thr * (w_i + w_j) / 2: each box may be shrunk by thr * w / 2 on both sides before the
// intersection test.  The reference's fp32 IoU is within ~1e-6 (relative) of the real-number value and its +1e-6 in the
// denominator only lowers it; 0.495 instead of 0.5 leaves 1 % of slack for that, and 1e-6 * max|coordinate| (8x the
// half-ulp of the shrunken coordinate) covers the rounding of the shrink itself.  Never negative; 0 = plain intersection.
__device__ __forceinline__ float nms_shrink_half(float lo, float hi, float thr01) {
    const float h = __fsub_rn(__fmul_rn(0.495f * thr01, __fsub_rn(hi, lo)), 1.0e-6f * fmaxf(fabsf(lo), fabsf(hi)));
    return h > 0.f ? h : 0.f;
}

// ---------------------------------------------------------------------------------------------------
// k_nms_suppress_zreg: k_nms_suppress_reg with the columns handed to the threads along a HILBERT CURVE over their centres.
// In k_nms_suppress_reg thread t owns the score ranks t, t+1024, ...: the 32 columns a warp holds in one slot are
// scattered over the image and every column meets every kept row of a tile (N^2/2 pair tests, ~8 instructions
// each).  Here the 1024 ranks of a slot are dealt out by a counting sort on the 8-bit Hilbert index of the box centre
// (16 x 16 grid over the image's own extent), so a warp's 32 columns of a slot are neighbours; the warp keeps the
// bounding box of its (shrunken, see nms_shrink_half) columns per slot, and per tile and slot ONE warp-wide test
// (two rows per lane, two ballots) picks the kept rows that can reach any of them -- the pair loop runs over those
// only (a few of 64 for COCO-sized boxes).  Which thread holds a column does not enter any verdict: every pair
// that can suppress is still decided by the same exact IoU, so keep lists are unchanged.
// Bookkeeping that followed the rank order moves to shared memory: the alive bitmap is indexed by rank (a dying
// column clears its bit with one atomicAnd; the tile's rows are read from it, and the final compaction too), the
// tile's boxes are read from the sorted list in global memory (1 KB per tile, L2), and s_perm[slot][thread] gives
// the rank a thread's column has inside the slot (needed only in the slot the current tile lies in).
// ---------------------------------------------------------------------------------------------------
// Hilbert index of cell (x, y) in a 16 x 16 grid (0..255): consecutive indices are always edge neighbours, so a run of
// cells never jumps across the image the way a Morton run does at quadrant borders.
__device__ __forceinline__ unsigned nms_hilbert16(int x, int y) {
    unsigned d = 0u;
#pragma unroll
    for (int s = 8; s > 0; s >>= 1) {
        const int rx = (x & s) ? 1 : 0, ry = (y & s) ? 1 : 0;
        d += (unsigned)(s * s * ((3 * rx) ^ ry));
        if (ry == 0) {
            if (rx == 1) { x = 15 - x; y = 15 - y; }
            const int t = x; x = y; y = t;
        }
    }
    return d;
}

template <bool CLASS_AWARE, int COLS>
__global__ void __launch_bounds__(kRegThreads, 1)
k_nms_suppress_zreg(const int *__restrict__ count, int A, const float4 *__restrict__ sbox, const float *__restrict__ sscore,
                    const int *__restrict__ scls, const int *__restrict__ sanchor, float thr, float img_size,
                    int *__restrict__ keep_anchor, float4 *__restrict__ keep_boxes, float *__restrict__ keep_scores,
                    int *__restrict__ keep_labels, int *__restrict__ keep_count, const int *__restrict__ only) {
    constexpr int kWarps = kRegThreads / 32;
    if (only && !only[blockIdx.x]) return;      // k_nms_suppress_grid has done this image
    __shared__ float4 s_tile[64];
    __shared__ float s_tarea[64];
    __shared__ int s_tcls[64];
    __shared__ unsigned long long s_diag[64];
    __shared__ float4 s_kbox[64];
    __shared__ float4 s_kshr[64];
    __shared__ float s_karea[64];
    __shared__ int s_kcls[64];
    __shared__ int s_nk;
    __shared__ unsigned s_alive[COLS * kWarps];                    // bit per score rank, 1 = alive
    __shared__ unsigned short s_perm[COLS * kRegThreads];          // [slot][thread] -> rank inside the slot, 0xFFFF = none
    __shared__ int s_hist[COLS][256];
    __shared__ float4 s_wbb[COLS][kWarps];                         // per slot and warp: bounding box of the warp's shrunken columns
    __shared__ float s_red[4][kWarps];
    const int b = blockIdx.x;
    const int n = count[b];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (n == 0) {
        if (tid == 0) keep_count[b] = 0;
        return;
    }
    const float4 *gbox = sbox + (size_t)b * A;
    const int *gcls = scls + (size_t)b * A;
    const float pinf = __int_as_float(0x7f800000), ninf = __int_as_float(0xff800000);
    const float4 never = make_float4(pinf, pinf, ninf, ninf);      // x2 - anything = -inf: never a candidate
    float4 bx[COLS];
    int cl[COLS];
    bool sane = true;      // finite, well-formed boxes: a pair that does not intersect then has IoU exactly 0
    // ---- pass 1 (rank order): sanity and the extent of the centres
    float mnx = pinf, mny = pinf, mxx = ninf, mxy = ninf;
#pragma unroll
    for (int k = 0; k < COLS; ++k) {
        const int c = k * kRegThreads + tid;
        bx[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < n) {
            bx[k] = gbox[c];
            const float sx = bx[k].x + bx[k].z, sy = bx[k].y + bx[k].w;      // twice the centre: only the order matters
            mnx = fminf(mnx, sx); mxx = fmaxf(mxx, sx);
            mny = fminf(mny, sy); mxy = fmaxf(mxy, sy);
        }
        sane = sane && (area_plain(bx[k]) >= 0.f) && (area_plain(bx[k]) < 1.0e38f) && (fabsf(bx[k].x) + fabsf(bx[k].y) + fabsf(bx[k].z) + fabsf(bx[k].w) < 3.0e38f);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
        mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    }
    if (lane == 0) { s_red[0][warp] = mnx; s_red[1][warp] = mny; s_red[2][warp] = mxx; s_red[3][warp] = mxy; }
    for (int i = tid; i < COLS * 256; i += kRegThreads) (&s_hist[0][0])[i] = 0;
    for (int i = tid; i < COLS * kRegThreads; i += kRegThreads) s_perm[i] = (unsigned short)0xFFFFu;
    for (int i = tid; i < COLS * kWarps; i += kRegThreads) {
        const int base = i * 32;
        s_alive[i] = n >= base + 32 ? 0xffffffffu : (n > base ? ((1u << (n - base)) - 1u) : 0u);
    }
    sane = __syncthreads_and(sane) && (0.f <= thr);
    const float thr01 = (thr >= 0.f && thr <= 1.f) ? thr : 0.f;      // outside [0,1] the candidate test is plain intersection
    mnx = s_red[0][lane]; mny = s_red[1][lane]; mxx = s_red[2][lane]; mxy = s_red[3][lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
        mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
    }
    const float gsx = (mxx > mnx && (mxx - mnx) < 3.0e38f) ? 16.0f / (mxx - mnx) : 0.f;
    const float gsy = (mxy > mny && (mxy - mny) < 3.0e38f) ? 16.0f / (mxy - mny) : 0.f;
    // ---- counting sort of every slot's 1024 ranks by Hilbert index (the order inside a bucket is whatever the atomics give:
    //      it decides which thread holds a column, nothing else)
    unsigned kp[COLS];
#pragma unroll
    for (int k = 0; k < COLS; ++k) {
        const int c = k * kRegThreads + tid;
        kp[k] = 0u;
        if (c < n) {
            const int kx = min(max((int)((bx[k].x + bx[k].z - mnx) * gsx), 0), 15);
            const int ky = min(max((int)((bx[k].y + bx[k].w - mny) * gsy), 0), 15);
            const unsigned key = nms_hilbert16(kx, ky);
            kp[k] = (key << 16) | (unsigned)atomicAdd(&s_hist[k][key], 1);
        }
    }
    __syncthreads();
    if (warp < COLS) {      // exclusive prefix over the slot's 256 buckets: 8 per lane + a warp scan
        int v[8], tot = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) { v[j] = s_hist[warp][lane * 8 + j]; tot += v[j]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        int run = inc - tot;
#pragma unroll
        for (int j = 0; j < 8; ++j) { s_hist[warp][lane * 8 + j] = run; run += v[j]; }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < COLS; ++k) {
        const int c = k * kRegThreads + tid;
        if (c < n) s_perm[k * kRegThreads + s_hist[k][kp[k] >> 16] + (int)(kp[k] & 0xFFFFu)] = (unsigned short)tid;
    }
    __syncthreads();
    // ---- pass 2: every thread takes the columns the sort dealt it; per slot the warp's bounding box of shrunken columns.
    //      Slot k is dealt ROTATED by 7*k warps: dealt straight, the nine slots of a warp would all lie at the same place
    //      along the curve, one region of the image, and a tile whose kept rows cluster there would load that warp nine
    //      times over while the others wait at the tile's barrier; rotated, a warp's slots sample nine regions.
    unsigned alive = 0u;
#pragma unroll
    for (int k = 0; k < COLS; ++k) {
        const unsigned r = s_perm[k * kRegThreads + ((tid + k * 7 * 32) & (kRegThreads - 1))];
        bx[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        cl[k] = -1;
        float4 sh = never;
        if (r != 0xFFFFu) {
            const int c = k * kRegThreads + (int)r;
            bx[k] = gbox[c];
            if (CLASS_AWARE) cl[k] = gcls[c];
            alive |= 1u << k;
            if (bx[k].z > bx[k].x && bx[k].w > bx[k].y) {
                const float hx = nms_shrink_half(bx[k].x, bx[k].z, thr01), hy = nms_shrink_half(bx[k].y, bx[k].w, thr01);
                sh = make_float4(__fadd_rn(bx[k].x, hx), __fadd_rn(bx[k].y, hy), __fsub_rn(bx[k].z, hx), __fsub_rn(bx[k].w, hy));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sh.x = fminf(sh.x, __shfl_xor_sync(0xffffffffu, sh.x, o)); sh.y = fminf(sh.y, __shfl_xor_sync(0xffffffffu, sh.y, o));
            sh.z = fmaxf(sh.z, __shfl_xor_sync(0xffffffffu, sh.z, o)); sh.w = fmaxf(sh.w, __shfl_xor_sync(0xffffffffu, sh.w, o));
        }
        if (lane == 0) s_wbb[k][warp] = sh;
    }
    __syncthreads();

    // the tile's rows come by rank from the sorted list in global memory; the next tile's are fetched before the column
    // pass of the current one, so their L2 latency is off the serial part of the tile
    float4 pre = make_float4(0.f, 0.f, 0.f, 0.f);
    int prec = -1;
    if (tid < 64 && tid < n) {
        pre = gbox[tid];
        if (CLASS_AWARE) prec = gcls[tid];
    }
    for (int t0 = 0; t0 < n; t0 += 64) {
        const int tn = min(64, n - t0);
        const int kt = t0 / kRegThreads;
        if (tid < 64) {
            s_tile[tid] = pre;
            s_tarea[tid] = area_plain(pre);
            if (CLASS_AWARE) s_tcls[tid] = prec;
        }
        __syncthreads();      // also: every column-pass write of the previous tile (s_alive) is done
        // -- diagonal 64x64 bitmask tile by warp ballots: warp w handles rows w and w + 32
        for (int r = warp; r < tn; r += kWarps) {
            const float4 bi = s_tile[r];
            const float ai = s_tarea[r];
            unsigned long long bits = 0ull;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int c = h * 32 + lane;
                bool sup = false;
                if (c > r && c < tn) {
                    sup = suppresses(bi, ai, s_tile[c], thr);
                    if (CLASS_AWARE) sup = sup && (s_tcls[c] == s_tcls[r]);
                }
                const unsigned m = __ballot_sync(0xffffffffu, sup);
                bits |= (unsigned long long)m << (32 * h);
            }
            if (lane == 0) s_diag[r] = bits;
        }
        __syncthreads();
        // -- greedy resolution of the tile by one warp ; the rows' alive bits come from, and
        //    go back to, the rank-indexed bitmap (nobody else touches the tile's two words now: column passes only
        //    clear bits of ranks >= t0 + 64)
        if (warp == 0) {
            const unsigned w0 = s_alive[t0 >> 5], w1 = tn > 32 ? s_alive[(t0 >> 5) + 1] : 0u;
            const unsigned long long d0 = lane < tn ? s_diag[lane] : 0ull;
            const unsigned long long d1 = lane + 32 < tn ? s_diag[lane + 32] : 0ull;
            unsigned long long rem = ~((unsigned long long)w0 | ((unsigned long long)w1 << 32));
            // only rows that suppress something inside the tile need a turn (in rank order); a row is kept iff its bit of
            // `rem` is clear at its turn, and later rows only set bits above themselves, so kept = ~rem at the end
            unsigned long long nz = (unsigned long long)__ballot_sync(0xffffffffu, d0 != 0ull) |
                                    ((unsigned long long)__ballot_sync(0xffffffffu, d1 != 0ull) << 32);
            while (nz) {
                const int r = __ffsll((long long)nz) - 1;
                nz &= nz - 1ull;
                const unsigned long long dr = __shfl_sync(0xffffffffu, r < 32 ? d0 : d1, r & 31);
                if (!((rem >> r) & 1ull)) rem |= dr;      // uniform across the warp
            }
            const unsigned long long kept = ~rem;
            const int nk = __popcll(kept);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int r = lane + 32 * h;
                if ((kept >> r) & 1ull) {
                    const int pos = __popcll(kept & ((1ull << r) - 1ull));
                    float4 v = s_tile[r];
                    float4 sh = v;
                    if (sane && !(v.z > v.x && v.w > v.y)) { v = never; sh = never; }
                    else {
                        const float hx = nms_shrink_half(v.x, v.z, thr01), hy = nms_shrink_half(v.y, v.w, thr01);
                        sh = make_float4(__fadd_rn(v.x, hx), __fadd_rn(v.y, hy), __fsub_rn(v.z, hx), __fsub_rn(v.w, hy));
                    }
                    s_kbox[pos] = v;
                    s_kshr[pos] = sh;
                    s_karea[pos] = s_tarea[r];
                    if (CLASS_AWARE) s_kcls[pos] = s_tcls[r];
                }
            }
            if (lane == 0) {
                s_nk = nk;
                s_alive[t0 >> 5] = (unsigned)(~rem);
                if (tn > 32) s_alive[(t0 >> 5) + 1] = (unsigned)((~rem) >> 32);
            }
        }
        __syncthreads();
        const int nk = s_nk;
        if (tid < 64) {
            pre = make_float4(0.f, 0.f, 0.f, 0.f);
            prec = -1;
            if (t0 + 64 + tid < n) {
                pre = gbox[t0 + 64 + tid];
                if (CLASS_AWARE) prec = gcls[t0 + 64 + tid];
            }
        }
        if (nk == 0) continue;
        // -- kept rows of this tile suppress every later, still-alive column
#pragma unroll
        for (int k = 0; k < COLS; ++k) {
            if (k < kt) continue;                                   // the whole slot ranks before the tile
            bool act = (alive >> k) & 1u;
            unsigned rk = 0u;
            if (k == kt) {                                          // the tile's own slot: only the later ranks
                rk = s_perm[k * kRegThreads + ((tid + k * 7 * 32) & (kRegThreads - 1))];
                act = act && (k * kRegThreads + (int)rk >= t0 + 64);
            }
            if (!__any_sync(0xffffffffu, act)) continue;
            const float4 bj = bx[k];
            bool dead = false;
            if (sane) {
                // which kept rows can reach any column of this warp's slot (>=, as the pair test counts a difference of +0 as a hit)
                const float4 bb = s_wbb[k][warp];
                const float4 r0 = s_kshr[lane & 63], r1 = s_kshr[(lane + 32) & 63];
                const bool h0 = lane < nk && r0.z >= bb.x && bb.z >= r0.x && r0.w >= bb.y && bb.w >= r0.y;
                const bool h1 = lane + 32 < nk && r1.z >= bb.x && bb.z >= r1.x && r1.w >= bb.y && bb.w >= r1.y;
                const unsigned long long rows = (unsigned long long)__ballot_sync(0xffffffffu, h0) |
                                                ((unsigned long long)__ballot_sync(0xffffffffu, h1) << 32);
                if (rows == 0ull) continue;
                if (act && bj.z > bj.x && bj.w > bj.y) {
                    float thr_k = thr01;
                    asm volatile("" : "+f"(thr_k));      // keeps the (tile-invariant) shrunken column out of registers across tiles
                    const float hx = nms_shrink_half(bj.x, bj.z, thr_k), hy = nms_shrink_half(bj.y, bj.w, thr_k);
                    const float jx = __fadd_rn(bj.x, hx), jy = __fadd_rn(bj.y, hy), jz = __fsub_rn(bj.z, hx), jw = __fsub_rn(bj.w, hy);
                    const float aj = area_plain(bj);
                    // the surviving rows, 32 at a time (32-bit find-first-set / clear-lowest: half the loop overhead of 64-bit)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        unsigned m = (unsigned)(rows >> (32 * h));
                        while (m && !dead) {
                            const int q = __ffs((int)m) - 1 + 32 * h;
                            m &= m - 1u;
                            const float4 bi = s_kshr[q];
                            // a > b  <=>  fl(a - b) > 0: four FADD sign bits ; +0 reads as a hit
                            const int sg = __float_as_int(__fsub_rn(bi.z, jx)) | __float_as_int(__fsub_rn(jz, bi.x)) |
                                           __float_as_int(__fsub_rn(bi.w, jy)) | __float_as_int(__fsub_rn(jw, bi.y));
                            if (sg >= 0) {
                                if (CLASS_AWARE && s_kcls[q] != cl[k]) continue;
                                // iou_nms, with the division taken only when the quotient could exceed thr: in the sane case
                                // (finite areas < 1e38, so nothing overflows) den > 0 and fl(inter / den) <= (inter / den) * (1 + 2^-24),
                                // hence inter <= 0.999 * thr * den  =>  IoU < thr.  Most candidates end here.
                                const float4 kb = s_kbox[q];
                                const float inter = inter_area(kb.x, kb.y, kb.z, kb.w, bj.x, bj.y, bj.z, bj.w);
                                const float den = YS_ADD(YS_SUB(YS_ADD(s_karea[q], aj), inter), kEpsNms);
                                if (inter > __fmul_rn(0.999f * thr, den)) dead = !(YS_DIV(inter, den) <= thr);
                            }
                        }
                    }
                }
            } else if (act) {
                const float aj = area_plain(bj);
                (void)aj;
                for (int q = 0; q < nk && !dead; ++q) {
                    if (CLASS_AWARE && s_kcls[q] != cl[k]) continue;
                    dead = suppresses(s_kbox[q], s_karea[q], bj, thr);
                }
            }
            if (dead) {
                alive &= ~(1u << k);
                if (k != kt) rk = s_perm[k * kRegThreads + ((tid + k * 7 * 32) & (kRegThreads - 1))];
                const int rank = k * kRegThreads + (int)rk;
                atomicAnd(&s_alive[rank >> 5], ~(1u << (rank & 31)));
            }
        }
        // no barrier here: the next tile's first barrier orders these reads of s_k* before warp 0 rewrites them
    }

    // -- survivors: the rank-indexed bitmap.  One warp turns the per-word popcounts into output offsets, then ALL warps
    //    write the kept detections (one 32-rank word per warp and trip; a single warp walking 8400 ranks with four
    //    dependent global loads per trip was 15 % of the kernel)
    __syncthreads();
    int *s_base = &s_hist[0][0];                                   // the sort's histogram is free now
    const int nwords = (n + 31) >> 5;
    if (warp == 0) {
        int running = 0;
        for (int w0 = 0; w0 < nwords; w0 += 32) {
            const int w = w0 + lane;
            const int cnt = w < nwords ? __popc(s_alive[w]) : 0;
            int inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            if (w < nwords) s_base[w] = running + inc - cnt;
            running += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) keep_count[b] = running;
    }
    __syncthreads();
    for (int w = warp; w < nwords; w += kWarps) {
        const unsigned m = s_alive[w];                             // bits of ranks >= n are zero
        if ((m >> lane) & 1u) {
            const size_t o = (size_t)b * A + s_base[w] + __popc(m & ((1u << lane) - 1u));
            const size_t s = (size_t)b * A + w * 32 + lane;
            if (keep_anchor) keep_anchor[o] = sanchor[s];
            if (keep_boxes) {
                const float4 v = sbox[s];
                // boxes / img_size, clamp to [0,1]  (train_yolo11_cuda.py:347-350)
                float4 q = make_float4(YS_DIV(v.x, img_size), YS_DIV(v.y, img_size), YS_DIV(v.z, img_size), YS_DIV(v.w, img_size));
                q.x = fminf2(fmaxf2(q.x, 0.f), 1.f); q.y = fminf2(fmaxf2(q.y, 0.f), 1.f);
                q.z = fminf2(fmaxf2(q.z, 0.f), 1.f); q.w = fminf2(fmaxf2(q.w, 0.f), 1.f);
                keep_boxes[o] = q;
            }
            if (keep_scores) keep_scores[o] = sscore[s];
            if (keep_labels) keep_labels[o] = scls[s];
        }
    }
}


// ---------------------------------------------------------------------------------------------------
// k_nms_suppress_grid: exact greedy NMS without the N^2 pair sweep.
// Greedy NMS keeps box j iff no KEPT box i of higher score has IoU(i, j) > thr (train_yolo11_cuda.py:361-399).  Two facts
// make that cheap: (1) IoU > thr forces the centres close -- inter > thr * union >= thr * area_a and inter_h <= h_a give
// inter_w > thr * w_a (same for b), and inter_w <= (w_a + w_b) / 2 - |dcx|, so |dcx| < (1 - thr) * max(w_a, w_b)
// (same in y): every possible suppressor of j lies within R = (1 - thr) * (largest box side of the image) of its centre;
// (2) the verdict on j only needs the verdicts on those few neighbours.  So: the image's sorted candidates are binned
// by centre into a 32 x 32 grid (counting sort in shared memory), and the keep flags are resolved in rounds: an
// undecided box scans the cells within R for higher-ranked boxes with IoU > thr (the reference's exact arithmetic) --
// one of them kept: removed; one of them undecided: wait for the next round; otherwise kept.  The lowest-ranked
// undecided box always decides, so the rounds terminate, and the fixed point is the sequential greedy result.
// With COCO-like boxes a scan meets ~1 % of the image's candidates and almost every box decides in the first round.
// One block per image, everything in shared memory.  Images this kernel does not take -- non-finite or inverted boxes,
// thr outside [0, 1), more candidates than fit, a search window wider than half the grid (one huge box), or a chain of
// dependencies longer than kGridMaxRounds -- are flagged in `fallback` and done by the tile kernels below.
// ---------------------------------------------------------------------------------------------------
constexpr int kGridThreads = 1024;
constexpr int kGridDim = 64;                 // cells per axis
constexpr int kGridCells = kGridDim * kGridDim;
constexpr int kGridMaxN = 8704;              // candidates per image held in shared memory
constexpr int kGridMaxRounds = 96;
#ifndef YS_NMS_GRID_MAX_WINDOW
#define YS_NMS_GRID_MAX_WINDOW 130
#endif
constexpr int kGridMaxWindow = YS_NMS_GRID_MAX_WINDOW;   // mean candidates per search window above which the tile kernel takes the image
constexpr int kCellsPerThread = kGridCells / kGridThreads;
static_assert(kGridCells % kGridThreads == 0, "the cell prefix sum gives every thread the same number of cells");
static_assert(kGridMaxN < (1 << 14), "rank and state share 16 bits");

// shared memory: boxes in cell order (16 B) | rank + state (2 B) | class (1 B, class-aware) | first position of every cell
// (+1) | two lists of undecided positions (the scatter cursors alias them) | kept bit per rank
static size_t grid_smem_bytes(int n_cap, bool class_aware) {
    const size_t lists = (size_t)2 * n_cap * 2, cursors = (size_t)kGridCells * 4;
    return (size_t)n_cap * (16 + 2 + (class_aware ? 1 : 0)) + 16 + (size_t)(kGridCells + 1) * 4 + 16 +
           (lists > cursors ? lists : cursors) + 16 + (size_t)((n_cap + 31) / 32) * 4 + 64;
}

template <bool CLASS_AWARE>
__global__ void __launch_bounds__(kGridThreads, 1)
k_nms_suppress_grid(const int *__restrict__ count, int A, const float4 *__restrict__ sbox, const float *__restrict__ sscore,
                    const int *__restrict__ scls, const int *__restrict__ sanchor, float thr, float img_size, int n_cap,
                    int *__restrict__ keep_anchor, float4 *__restrict__ keep_boxes, float *__restrict__ keep_scores,
                    int *__restrict__ keep_labels, int *__restrict__ keep_count, int *__restrict__ fallback) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    constexpr int kWarps = kGridThreads / 32;
    constexpr unsigned kRankMask = 0x3FFFu, kKept = 1u << 14, kRemoved = 2u << 14;      // state in the two top bits of s_rs
    auto align16 = [](void *p) { return reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(p) + 15) & ~(uintptr_t)15); };
    float4 *s_box = reinterpret_cast<float4 *>(s_raw);                                  // [n_cap] in cell order
    unsigned short *s_rs = reinterpret_cast<unsigned short *>(s_box + n_cap);           // [n_cap] score rank | state << 14
    unsigned char *s_cls = reinterpret_cast<unsigned char *>(s_rs + n_cap);             // [n_cap] (class-aware only)
    int *s_cell = reinterpret_cast<int *>(align16(s_cls + (CLASS_AWARE ? n_cap : 0)));  // [cells + 1] first position of a cell
    int *s_cursor = reinterpret_cast<int *>(align16(s_cell + kGridCells + 1));          // [cells] histogram / scatter cursors ...
    unsigned short *s_list0 = reinterpret_cast<unsigned short *>(s_cursor);             // ... then: undecided positions, two buffers
    unsigned short *s_list1 = s_list0 + n_cap;
    const size_t lists = (size_t)2 * n_cap * 2, cursors = (size_t)kGridCells * 4;
    unsigned *s_alive = reinterpret_cast<unsigned *>(align16(reinterpret_cast<unsigned char *>(s_cursor) + (lists > cursors ? lists : cursors)));
    __shared__ float s_red[7][kWarps];
    __shared__ int s_part[kWarps];
    __shared__ int s_nlist[2];
    constexpr int kWorkClasses = 128;
    __shared__ int s_hist[kWorkClasses];
    const int b = blockIdx.x;
    const int n = count[b];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (n == 0) {
        if (tid == 0) { keep_count[b] = 0; fallback[b] = 0; }
        return;
    }
    const float4 *gbox = sbox + (size_t)b * A;
    const int *gcls = scls + (size_t)b * A;
    const float pinf = __int_as_float(0x7f800000), ninf = __int_as_float(0xff800000);
    // ---- pass 1 (rank order): sanity, extent of the (doubled) centres, largest width / height, largest |coordinate|
    bool sane = n <= n_cap && thr >= 0.f && thr < 1.f;
    float mnx = pinf, mny = pinf, mxx = ninf, mxy = ninf, wmax = 0.f, hmax = 0.f, big = 0.f;
    if (sane) {
        for (int r = tid; r < n; r += kGridThreads) {
            const float4 v = gbox[r];
            const float w = v.z - v.x, h = v.w - v.y;
            sane = sane && (w >= 0.f) && (h >= 0.f) && (fabsf(v.x) + fabsf(v.y) + fabsf(v.z) + fabsf(v.w) < 1.0e18f);
            if (CLASS_AWARE) sane = sane && ((unsigned)gcls[r] < 256u);
            const float sx = v.x + v.z, sy = v.y + v.w;
            mnx = fminf(mnx, sx); mxx = fmaxf(mxx, sx);
            mny = fminf(mny, sy); mxy = fmaxf(mxy, sy);
            wmax = fmaxf(wmax, w); hmax = fmaxf(hmax, h);
            big = fmaxf(big, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
        }
    }
    auto wred = [&]() {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
            mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o)); mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
            wmax = fmaxf(wmax, __shfl_xor_sync(0xffffffffu, wmax, o)); hmax = fmaxf(hmax, __shfl_xor_sync(0xffffffffu, hmax, o));
            big = fmaxf(big, __shfl_xor_sync(0xffffffffu, big, o));
        }
    };
    wred();
    if (lane == 0) { s_red[0][warp] = mnx; s_red[1][warp] = mny; s_red[2][warp] = mxx; s_red[3][warp] = mxy; s_red[4][warp] = wmax; s_red[5][warp] = hmax; s_red[6][warp] = big; }
    for (int i = tid; i < kGridCells; i += kGridThreads) s_cursor[i] = 0;
    sane = __syncthreads_and(sane);
    mnx = s_red[0][lane]; mny = s_red[1][lane]; mxx = s_red[2][lane]; mxy = s_red[3][lane]; wmax = s_red[4][lane]; hmax = s_red[5][lane]; big = s_red[6][lane];
    wred();
    // Search radii in doubled-centre units (see the note above the kernel): per axis min((1 - thr) * largest side of the image,
    // (1 - thr) / thr * own side) -- IoU > thr also bounds the partner's side by own side / thr -- with 1 % of slack for the
    // reference's fp32 rounding of the IoU (its +1e-6 in the denominator only lowers it) and an absolute term for the
    // rounding of the centre sums.
    const float kf = 1.0f - thr + 0.01f, slack = 4.0e-6f * big;
    const float kown = thr > 0.f ? 2.0f * kf / thr : pinf;
    const float R2xg = 2.0f * kf * wmax + slack, R2yg = 2.0f * kf * hmax + slack;
    const float gsx = (mxx > mnx) ? (float)kGridDim / (mxx - mnx) : 0.f;
    const float gsy = (mxy > mny) ? (float)kGridDim / (mxy - mny) : 0.f;
    // a window wider than half the grid (one huge box, or everything in one spot) is no pruning at all
    const bool wide = (gsx > 0.f && R2xg * gsx > 0.25f * kGridDim) || (gsy > 0.f && R2yg * gsy > 0.25f * kGridDim) ||
                      (gsx == 0.f && gsy == 0.f && n > 256);
    if (!sane || wide) {              // block-uniform
        if (tid == 0) fallback[b] = 1;
        return;
    }
    auto cell_x = [&](float c2) { return min(max((int)((c2 - mnx) * gsx), 0), kGridDim - 1); };
    auto cell_y = [&](float c2) { return min(max((int)((c2 - mny) * gsy), 0), kGridDim - 1); };
    // ---- counting sort by cell
    for (int r = tid; r < n; r += kGridThreads) {
        const float4 v = gbox[r];
        atomicAdd(&s_cursor[cell_y(v.y + v.w) * kGridDim + cell_x(v.x + v.z)], 1);
    }
    __syncthreads();
    {   // exclusive prefix over the cells: kCellsPerThread consecutive cells per thread
        int v[kCellsPerThread], tot = 0;
#pragma unroll
        for (int j = 0; j < kCellsPerThread; ++j) { v[j] = s_cursor[tid * kCellsPerThread + j]; tot += v[j]; }
        int inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_part[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const int pv = s_part[lane];
            int pi = pv;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, pi, o);
                if (lane >= o) pi += t;
            }
            s_part[lane] = pi - pv;
        }
        __syncthreads();
        int run = s_part[warp] + inc - tot;
#pragma unroll
        for (int j = 0; j < kCellsPerThread; ++j) {
            s_cell[tid * kCellsPerThread + j] = run;
            s_cursor[tid * kCellsPerThread + j] = run;
            run += v[j];
        }
        if (tid == kGridThreads - 1) s_cell[kGridCells] = run;
    }
    __syncthreads();
    for (int r = tid; r < n; r += kGridThreads) {
        const float4 v = gbox[r];
        const int p = atomicAdd(&s_cursor[cell_y(v.y + v.w) * kGridDim + cell_x(v.x + v.z)], 1);
        s_box[p] = v;
        s_rs[p] = (unsigned short)r;                 // state 0: undecided
        if (CLASS_AWARE) s_cls[p] = (unsigned char)gcls[r];
    }
    if (tid == 0) { s_nlist[0] = 0; s_nlist[1] = 0; s_part[0] = 0; }
    __syncthreads();                                  // the cursors are dead from here: their memory holds the lists
    // ---- is the image sparse enough?  The scan below visits every candidate of a box's window; dense scenes (boxes jittered
    //      around a few centres, low thresholds) have hundreds of candidates per window and are better served by the tile
    //      kernel, whose cost does not depend on the density.  The window populations come from the cell table for free.
    //      The same pass files every box under a work class (its window population + a charge per row), and round 0 then
    //      walks the boxes in order of descending class: the 32 lanes of a warp get windows of similar size.  (Walked in
    //      cell order, a warp's lanes had windows between a few and a few hundred candidates: ncu read 11 of 32 lanes
    //      active per instruction.)  Results do not depend on the order the boxes of a round are visited in.
    {
        for (int i = tid; i < kWorkClasses; i += kGridThreads) s_hist[i] = 0;
        __syncthreads();
        long long pop = 0;
        for (int p = tid; p < n; p += kGridThreads) {
            const float4 a = s_box[p];
            const float c2x = a.x + a.z, c2y = a.y + a.w;
            const float R2x = fminf(R2xg, kown * (a.z - a.x) + slack), R2y = fminf(R2yg, kown * (a.w - a.y) + slack);
            const int cx0 = cell_x(c2x - R2x), cx1 = cell_x(c2x + R2x), cy0 = cell_y(c2y - R2y), cy1 = cell_y(c2y + R2y);
            int mine = 0;
            for (int cy = cy0; cy <= cy1; ++cy) mine += s_cell[cy * kGridDim + cx1 + 1] - s_cell[cy * kGridDim + cx0];
            pop += mine;
            const int wc = kWorkClasses - 1 - min((mine + 2 * (cy1 - cy0 + 1)) >> 2, kWorkClasses - 1);    // descending work
            s_list0[p] = (unsigned short)wc;
            atomicAdd(&s_hist[wc], 1);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pop += __shfl_xor_sync(0xffffffffu, pop, o);
        if (lane == 0) atomicAdd(&s_part[0], (int)min(pop, (long long)0x3fffff));
        __syncthreads();
        if ((long long)s_part[0] > (long long)kGridMaxWindow * n) {      // block-uniform
            if (tid == 0) fallback[b] = 1;
            return;
        }
        if (warp == 0) {                                   // exclusive prefix over the work classes
            int run = 0;
            for (int c0 = 0; c0 < kWorkClasses; c0 += 32) {
                const int v = s_hist[c0 + lane];
                int inc = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += t;
                }
                s_hist[c0 + lane] = run + inc - v;
                run += __shfl_sync(0xffffffffu, inc, 31);
            }
        }
        __syncthreads();
        for (int p = tid; p < n; p += kGridThreads) s_list1[atomicAdd(&s_hist[s_list0[p]], 1)] = (unsigned short)p;
        __syncthreads();
    }
    // ---- rounds.  Round 0 walks every position; later rounds only the positions still undecided (compacted lists).
    volatile unsigned short *rs = s_rs;
    int rounds = 0;
    bool gave_up = false;
    while (true) {
        const unsigned short *lin = (rounds & 1) ? s_list0 : s_list1;        // written by the previous round
        unsigned short *lout = (rounds & 1) ? s_list1 : s_list0;
        int *nout = &s_nlist[rounds & 1 ? 1 : 0];
        const int nin = rounds == 0 ? n : s_nlist[rounds & 1 ? 0 : 1];      // round 0: s_list1 = all positions, heaviest first
        for (int i = tid; i < nin; i += kGridThreads) {
            const int p = (int)lin[i];
            const float4 a = s_box[p];
            const unsigned ra = rs[p] & kRankMask;
            const float c2x = a.x + a.z, c2y = a.y + a.w;
            const float R2x = fminf(R2xg, kown * (a.z - a.x) + slack), R2y = fminf(R2yg, kown * (a.w - a.y) + slack);
            const int cx0 = cell_x(c2x - R2x), cx1 = cell_x(c2x + R2x), cy0 = cell_y(c2y - R2y), cy1 = cell_y(c2y + R2y);
            const float area_a = area_plain(a);
            const float tw_a = fmaxf(0.99f * thr * (a.z - a.x), 0.f), th_a = fmaxf(0.99f * thr * (a.w - a.y), 0.f);
            const int cls_a = CLASS_AWARE ? (int)s_cls[p] : 0;
            const float thr999 = 0.999f * thr;
            bool removed = false, pending = false;
            // one flat loop over the window's rows: a lane moves on to its next row without waiting for the other lanes' rows
            // (the step to the next row is predicated, not a branch: with ~10 candidates per row some lane of the warp changes
            // rows in almost every trip; an empty row costs its lane one idle trip.  Loading the next trip's candidate under
            // this trip's arithmetic -- software pipelining -- was measured: +17 % instructions, 764 -> 796 us, not taken)
            int cy = cy0, q = s_cell[cy0 * kGridDim + cx0], q1 = s_cell[cy0 * kGridDim + cx1 + 1];
            {
                for (;;) {
                    const bool adv = q >= q1;
                    if (adv && cy >= cy1) break;
                    cy += adv ? 1 : 0;
                    const int cb = cy * kGridDim;
                    const int nq = s_cell[cb + cx0], nq1 = s_cell[cb + cx1 + 1];
                    q = adv ? nq : q;
                    q1 = adv ? nq1 : q1;
                    const bool valid = q < q1;
                    const int qq = q;
                    q += valid ? 1 : 0;
                    // straight-line candidate test (no early `continue`s: a warp's lanes walk different candidates, and
                    // every divergent exit was being executed by the whole warp): higher-scored, not removed, same class,
                    // and -- IoU > thr needs inter_w > thr * own width and inter_h > thr * own height (see the note above
                    // the kernel; 1 % of slack for the reference's fp32 rounding) -- two compares that settle almost every pair
                    const unsigned rq = rs[qq];                  // qq <= n: an idle trip reads inside the shared arrays, its result is masked
                    const float4 e = s_box[qq];
                    const float iw = YS_SUB(fminf2(e.z, a.z), fmaxf2(e.x, a.x));
                    const float ih = YS_SUB(fminf2(e.w, a.w), fmaxf2(e.y, a.y));
                    bool cand = valid & ((rq & kRankMask) < ra) & !(rq & kRemoved) & (iw > tw_a) & (ih > th_a);
                    if (CLASS_AWARE) cand = cand & ((int)s_cls[qq] == cls_a);
                    // iou_nms(e, a) > thr with the reference's arithmetic (train_yolo11_cuda.py:418-435, :396).  inter and den
                    // are computed for every candidate (a warp's lanes walk different candidates: ~12 % pass the gate, so a
                    // branch here was taken by almost every warp with a handful of lanes); the division only when the
                    // quotient could exceed thr (den > 0; fl(x / y) <= (x / y) * (1 + 2^-24))
                    const float inter = YS_MUL(iw, ih);
                    const float den = YS_ADD(YS_SUB(YS_ADD(area_plain(e), area_a), inter), kEpsNms);
                    if (cand & (inter > __fmul_rn(thr999, den))) {
                        if (!(YS_DIV(inter, den) <= thr)) {
                            if (rq & kKept) { removed = true; break; }
                            pending = true;                                // undecided when read (it may have been decided
                        }                                                  // since: then the next round sees it)
                    }
                }
            }
            if (removed) rs[p] = (unsigned short)(ra | kRemoved);
            else if (!pending) rs[p] = (unsigned short)(ra | kKept);
            else lout[atomicAdd(nout, 1)] = (unsigned short)p;
        }
        __syncthreads();
        const int left = *nout;
        if (left == 0) break;
        if (++rounds >= kGridMaxRounds) { gave_up = true; break; }
        if (tid == 0) s_nlist[rounds & 1 ? 1 : 0] = 0;                    // the list this coming round appends to
        __syncthreads();
    }
#ifdef YS_NMS_GRID_DEBUG
    if (tid == 0 && b < 6) printf("GRID b=%d n=%d rounds=%d gave_up=%d R2xg=%.1f cells=%.2f left=%d\n", b, n, rounds, (int)gave_up, R2xg, R2xg * gsx, *(&s_nlist[rounds & 1 ? 1 : 0]));
#endif
    if (gave_up) {                    // block-uniform: a long dependency chain; the tile kernel redoes this image
        if (tid == 0) fallback[b] = 1;
        return;
    }
    // ---- survivors: rank-indexed bitmap, per-word offsets from one popcount scan, all warps write
    for (int i = tid; i < (n + 31) / 32; i += kGridThreads) s_alive[i] = 0u;
    __syncthreads();
    for (int p = tid; p < n; p += kGridThreads) {
        const unsigned v = s_rs[p];
        if (v & kKept) atomicOr(&s_alive[(v & kRankMask) >> 5], 1u << (v & 31u));
    }
    __syncthreads();
    int *s_base = s_cursor;                                        // the lists are free now
    const int nwords = (n + 31) >> 5;
    if (warp == 0) {
        int running = 0;
        for (int w0 = 0; w0 < nwords; w0 += 32) {
            const int w = w0 + lane;
            const int cnt = w < nwords ? __popc(s_alive[w]) : 0;
            int inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            if (w < nwords) s_base[w] = running + inc - cnt;
            running += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) { keep_count[b] = running; fallback[b] = 0; }
    }
    __syncthreads();
    for (int w = warp; w < nwords; w += kWarps) {
        const unsigned m = s_alive[w];
        if ((m >> lane) & 1u) {
            const size_t o = (size_t)b * A + s_base[w] + __popc(m & ((1u << lane) - 1u));
            const size_t src = (size_t)b * A + w * 32 + lane;
            if (keep_anchor) keep_anchor[o] = sanchor[src];
            if (keep_boxes) {
                const float4 v = sbox[src];
                // boxes / img_size, clamp to [0,1]  (train_yolo11_cuda.py:347-350)
                float4 q = make_float4(YS_DIV(v.x, img_size), YS_DIV(v.y, img_size), YS_DIV(v.z, img_size), YS_DIV(v.w, img_size));
                q.x = fminf2(fmaxf2(q.x, 0.f), 1.f); q.y = fminf2(fmaxf2(q.y, 0.f), 1.f);
                q.z = fminf2(fmaxf2(q.z, 0.f), 1.f); q.w = fminf2(fmaxf2(q.w, 0.f), 1.f);
                keep_boxes[o] = q;
            }
            if (keep_scores) keep_scores[o] = sscore[src];
            if (keep_labels) keep_labels[o] = scls[src];
        }
    }
}

struct NmsWorkspace {
    unsigned long long *keys, *keys2;      // keys2: segment-sorted keys, only when A > kSortMaxN
    float4 *boxes, *sbox;
    int *cls, *scls, *sanchor, *count, *fallback;
    float *sscore;
    size_t bytes;
};
static void carve_nms(NmsWorkspace &W, void *base, size_t cap, int B, int A) {
    Carver c(base, cap);
    const size_t n = (size_t)B * A;
    W.count = c.take<int>((size_t)B + 1);
    W.fallback = c.take<int>((size_t)B + 1);
    W.keys = c.take<unsigned long long>(n);
    W.keys2 = A > kSortMaxN ? c.take<unsigned long long>(n) : nullptr;
    W.boxes = c.take<float4>(n);
    W.sbox = c.take<float4>(n);
    W.cls = c.take<int>(n);
    W.scls = c.take<int>(n);
    W.sanchor = c.take<int>(n);
    W.sscore = c.take<float>(n);
    W.bytes = align_up(c.off, 256);
}

static size_t suppress_smem_bytes(int A, int smem_boxes) {
    const int max_words = (A + 31) / 32;
    return (size_t)smem_boxes * sizeof(float4) + (size_t)((max_words + 3) & ~3) * 4 + 64 * 8 + 64 * 16 + 64 * 4 + 64 * 4 + 65 * 4 + 64;
}

static int run_sort_suppress(const NmsWorkspace &W, int B, int A, float iou_thr, float img_size, int class_aware,
                             const float4 *boxes, const int *cls, int32_t *keep_anchor, float *keep_boxes,
                             float *keep_scores, int32_t *keep_labels, int32_t *keep_count, cudaStream_t stream) {
    int P = 1;
    while (P < A && P < kSortMaxN) P <<= 1;
    // k_nms_sort needs H + P2 slots (its two runs), at most Hm + pow2ceil(A - Hm) over all counts n <= A, Hm = the largest
    // power of two <= A: 8448 slots = 66 KB at A = 8400, so two blocks share an SM (256 images: one wave instead of two)
    size_t sort_slots = (size_t)P;
    if (A <= kSortMaxN) {
        int Hm = 1;
        while (Hm * 2 <= A) Hm <<= 1;
        int P2m = 0;
        if (A > Hm) { P2m = 1; while (P2m < A - Hm) P2m <<= 1; }
        sort_slots = (size_t)Hm + P2m;
    }
    const size_t sort_smem = sort_slots * sizeof(unsigned long long);
    static DeviceOnce once_sort, once_sup[2];
    int rc = optin_smem((const void *)k_nms_sort, kSortMaxN * 8, once_sort, "k_nms_sort");
    if (rc) return rc;
    rc = optin_smem((const void *)k_nms_suppress<false>, 220 * 1024, once_sup[0], "k_nms_suppress");
    if (rc) return rc;
    rc = optin_smem((const void *)k_nms_suppress<true>, 220 * 1024, once_sup[1], "k_nms_suppress");
    if (rc) return rc;
    if (A <= kSortMaxN) {
        k_nms_sort<<<B, kSortThreads, sort_smem, stream>>>(W.keys, W.count, A, boxes, cls, W.sbox, W.sscore, W.scls, W.sanchor);
        YS_CHECK_LAUNCH("k_nms_sort");
    } else {
        static DeviceOnce once_seg;
        rc = optin_smem((const void *)k_nms_sort_seg, kSortMaxN * 8, once_seg, "k_nms_sort_seg");
        if (rc) return rc;
        YS_REQUIRE(B <= 65535, YS_ERR_UNSUPPORTED, "B must be <= 65535 (grid.y) when A > %d", kSortMaxN);
        const int nseg = (A + kSortMaxN - 1) / kSortMaxN;
        k_nms_sort_seg<<<dim3(nseg, B), kSortThreads, (size_t)kSortMaxN * 8, stream>>>(W.keys, W.count, A, W.keys2);
        YS_CHECK_LAUNCH("k_nms_sort_seg");
        k_nms_merge<<<dim3((A + 255) / 256, B), 256, 0, stream>>>(W.keys2, W.count, A, boxes, cls, W.sbox, W.sscore, W.scls, W.sanchor);
        YS_CHECK_LAUNCH("k_nms_merge");
    }
    // Grid kernel first (every image it can take); the tile kernels then run for the images it flagged only.
    const int *only = nullptr;
#ifndef YS_NMS_NO_GRID
    if (A <= kGridMaxN) {
        static DeviceOnce once_grid[2];
        const int n_cap = (A + 31) & ~31;
        const size_t gsm = grid_smem_bytes(n_cap, class_aware != 0);
        rc = class_aware ? optin_smem((const void *)k_nms_suppress_grid<true>, grid_smem_bytes(kGridMaxN, true), once_grid[1], "k_nms_suppress_grid")
                         : optin_smem((const void *)k_nms_suppress_grid<false>, grid_smem_bytes(kGridMaxN, false), once_grid[0], "k_nms_suppress_grid");
        if (rc) return rc;
        if (class_aware)
            k_nms_suppress_grid<true><<<B, kGridThreads, gsm, stream>>>(W.count, A, W.sbox, W.sscore, W.scls, W.sanchor, iou_thr, img_size, n_cap,
                                                                         keep_anchor, (float4 *)keep_boxes, keep_scores, keep_labels, keep_count, W.fallback);
        else
            k_nms_suppress_grid<false><<<B, kGridThreads, gsm, stream>>>(W.count, A, W.sbox, W.sscore, W.scls, W.sanchor, iou_thr, img_size, n_cap,
                                                                          keep_anchor, (float4 *)keep_boxes, keep_scores, keep_labels, keep_count, W.fallback);
        YS_CHECK_LAUNCH("k_nms_suppress_grid");
        only = W.fallback;
    }
#endif
    if (A <= 9 * kRegThreads) {      // register-resident columns (every 640 px / 8400-candidate case)
#define YS_LAUNCH_REG(CA, COLS)                                                                                              \
    k_nms_suppress_zreg<CA, COLS><<<B, kRegThreads, 0, stream>>>(W.count, A, W.sbox, W.sscore, W.scls, W.sanchor, iou_thr,      \
                                                                img_size, keep_anchor, (float4 *)keep_boxes, keep_scores,     \
                                                                keep_labels, keep_count, only)
        if (A <= 2 * kRegThreads) { if (class_aware) YS_LAUNCH_REG(true, 2); else YS_LAUNCH_REG(false, 2); }
        else if (A <= 5 * kRegThreads) { if (class_aware) YS_LAUNCH_REG(true, 5); else YS_LAUNCH_REG(false, 5); }
        else { if (class_aware) YS_LAUNCH_REG(true, 9); else YS_LAUNCH_REG(false, 9); }
#undef YS_LAUNCH_REG
        YS_CHECK_LAUNCH("k_nms_suppress_zreg");
        return YS_OK;
    }
    const int smem_boxes = A <= kSupSmemBoxes ? A : 0;
    const size_t sup_smem = suppress_smem_bytes(A, smem_boxes);
    if (class_aware)
        k_nms_suppress<true><<<B, kSupThreads, sup_smem, stream>>>(W.count, A, W.sbox, W.sscore, W.scls, W.sanchor, iou_thr, img_size,
                                                                   smem_boxes, keep_anchor, (float4 *)keep_boxes, keep_scores, keep_labels, keep_count, only);
    else
        k_nms_suppress<false><<<B, kSupThreads, sup_smem, stream>>>(W.count, A, W.sbox, W.sscore, W.scls, W.sanchor, iou_thr, img_size,
                                                                    smem_boxes, keep_anchor, (float4 *)keep_boxes, keep_scores, keep_labels, keep_count, only);
    YS_CHECK_LAUNCH("k_nms_suppress");
    return YS_OK;
}

}  // namespace ys

using namespace ys;

extern "C" size_t ys_nms_workspace_bytes(int B, int A) {
    NmsWorkspace W;
    carve_nms(W, nullptr, 0, B > 0 ? B : 1, A > 0 ? A : 1);
    return W.bytes;
}

extern "C" int ys_nms_batched(const float *pred, int layout, int B, int A, int nc, float conf_thr, float iou_thr,
                              float img_size, int class_aware, int32_t *keep_anchor, float *keep_boxes,
                              float *keep_scores, int32_t *keep_labels, int32_t *keep_count, void *workspace,
                              size_t workspace_bytes, void *stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    YS_REQUIRE(pred && keep_count && workspace, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE(B >= 1 && A >= 1 && nc >= 1, YS_ERR_BAD_ARG, "B, A, nc must be >= 1");
    YS_REQUIRE(layout == 0 || layout == 1, YS_ERR_BAD_ARG, "layout must be 0 (B,4+nc,A) or 1 (B,A,4+nc)");
    YS_REQUIRE(!keep_boxes || ((uintptr_t)keep_boxes & 15) == 0, YS_ERR_BAD_ARG, "keep_boxes must be 16-byte aligned");
    NmsWorkspace W;
    carve_nms(W, workspace, workspace_bytes, B, A);
    YS_REQUIRE(W.bytes <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu", W.bytes, workspace_bytes);
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    cudaMemsetAsync(W.count, 0, sizeof(int) * ((size_t)B + 1), stream);
    const long long warps = (long long)B * ((A + 31) / 32);
    const unsigned blocks = (unsigned)((warps + kFilterThreads / 32 - 1) / (kFilterThreads / 32));
    if (layout == 1) {
        const size_t smem = (size_t)(kFilterThreads / 32) * 32 * ((4 + nc) | 1) * sizeof(float);
        YS_REQUIRE(smem <= 200 * 1024, YS_ERR_UNSUPPORTED, "nc too large for the staged filter (%d)", nc);
        static DeviceOnce once_filter;
        const int rcf = optin_smem((const void *)k_nms_filter<1>, 200 * 1024, once_filter, "k_nms_filter");
        if (rcf) return rcf;
        k_nms_filter<1><<<blocks, kFilterThreads, smem, stream>>>(pred, B, A, nc, conf_thr, W.keys, W.boxes, W.cls, W.count);
    } else {
        k_nms_filter<0><<<blocks, kFilterThreads, 0, stream>>>(pred, B, A, nc, conf_thr, W.keys, W.boxes, W.cls, W.count);
    }
    YS_CHECK_LAUNCH("k_nms_filter");
    return run_sort_suppress(W, B, A, iou_thr, img_size, class_aware, W.boxes, W.cls, keep_anchor, keep_boxes, keep_scores,
                             keep_labels, keep_count, stream);
}

extern "C" int ys_nms_single(const float *boxes, const float *scores, const int32_t *labels, int N, float iou_thr,
                             int32_t *keep, int32_t *keep_count, void *workspace, size_t workspace_bytes, void *stream_v) {
    cudaStream_t stream = (cudaStream_t)stream_v;
    YS_REQUIRE(boxes && scores && keep && keep_count && workspace, YS_ERR_BAD_ARG, "null pointer");
    YS_REQUIRE(N >= 1, YS_ERR_BAD_ARG, "N must be >= 1");
    YS_REQUIRE(((uintptr_t)boxes & 15) == 0, YS_ERR_BAD_ARG, "boxes must be 16-byte aligned");
    NmsWorkspace W;
    carve_nms(W, workspace, workspace_bytes, 1, N);
    YS_REQUIRE(W.bytes <= workspace_bytes, YS_ERR_WORKSPACE, "workspace too small: need %zu, got %zu", W.bytes, workspace_bytes);
    YS_REQUIRE(((uintptr_t)workspace & 255) == 0, YS_ERR_BAD_ARG, "workspace must be 256-byte aligned");
    k_nms_keys_single<<<(N + 255) / 256, 256, 0, stream>>>(scores, N, W.keys, W.count);
    YS_CHECK_LAUNCH("k_nms_keys_single");
    return run_sort_suppress(W, 1, N, iou_thr, 1.0f, labels != nullptr, (const float4 *)boxes, labels, keep, nullptr, nullptr,
                             nullptr, keep_count, stream);
}
