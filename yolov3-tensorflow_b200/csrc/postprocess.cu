// Score thresholding and class-aware greedy NMS (yolov3/yolov3_post_process.py).
//
//   filter  _filter_single_head_boxes :44-76   score = max_c p_c * conf (fp32), class = FIRST arg-max,
//           keep iff score > fp32(thresh) (strict); rows [x0/W, y0/H, x1/W, y1/H, conf, prob, cls, score]
//           in head /8,/16,/32 order, raster (y, x, b) inside a head.
//   nms     _apply_nms :108-129, _cal_iou :132-145   stable sort by score descending, a later box is
//           dropped iff an earlier KEPT box of the same class has IoU_f64 > thresh (strict).
//
// Kernels:
//   score_packed_kernel   one sweep over the packed logits (HBM-bound, 4*A*L bytes/image): tile in shared
//                         memory by bulk copy, one thread per record, (score, prob, class) written in
//                         head-raster order (12 B/record).
//   compact_kernel        one CTA per image: order-preserving compaction (block scan per head) of the
//                         records that pass the threshold; survivors' boxes are re-decoded from 5 logits.
//   sort_kernel           one CTA per image: 64-bit keys (class | ~score | row) bitonic-sorted in shared
//                         memory (global memory beyond 16 Ki candidates).
//   nms_kernel            one CTA per (class, image) segment: 256-candidate chunks; each chunk is first
//                         tested against the boxes kept so far, then resolved internally with a
//                         shared-memory IoU bit matrix and a warp sweep.  IoU is fp64, uncontracted, in the
//                         reference's operation order.
#include <string.h>

#include "tile_ops.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kScoreThreads = 128;

// ---- (score, prob, class) of every record -----------------------------------------------------------
__global__ void __launch_bounds__(kScoreThreads) score_packed_kernel(const __grid_constant__ Geo g,
                                                                     const float* __restrict__ predicts,
                                                                     const TileLaunch q, float* __restrict__ score,
                                                                     float* __restrict__ prob, int32_t* __restrict__ cls) {
    extern __shared__ __align__(128) unsigned char ssm[];
    __shared__ uint64_t bar;
    const int L = g.L, C = g.C, tr = q.tr;
    float* tile = (float*)ssm;
    const int tid = threadIdx.x;
    const long long rec0 = (long long)blockIdx.x * tr;
    const int nrec = (int)min((long long)tr, q.total_rec - rec0);
    tile_load<kScoreThreads>(tile, predicts + (size_t)rec0 * L, nrec * L, q.tma, &bar);

    const int off1 = g.head_rec[0], off2 = g.head_rec[0] + g.head_rec[1];
    // one thread per record: records are L floats apart in shared memory (bank-conflict free for odd L)
    for (int ri = tid; ri < nrec; ri += kScoreThreads) {
        const long long grec = rec0 + ri;
        const int n_img = (int)(grec / g.rpi);
        const RecLoc p = locate(g, (int)(grec - (long long)n_img * g.rpi));
        const int rih = (p.y * g.W[p.h] + p.x) * g.B[p.h] + p.b;
        const size_t o = (size_t)n_img * g.rpi + (p.h == 0 ? 0 : (p.h == 1 ? off1 : off2)) + rih;
        const float* r = tile + (size_t)ri * L;
        const float conf = clip_ref(sigmoid_ref(r[4]), g.lo, g.hi);
        if (C > 0) {
            const ClassStats cs = class_stats(r + 5, C);
            float best;
            int arg;
            class_best(r + 5, C, cs, g.lo, g.hi, best, arg);
            score[o] = __fmul_rn(best, conf);                           // :59
            prob[o] = best;
            cls[o] = arg;
        } else {
            score[o] = conf;                                            // :53-55 (prob 1, class 0)
            prob[o] = 1.0f;
            cls[o] = 0;
        }
    }
}

// Same outputs from already decoded heads of ONE image (filter_boxes' own inputs).
__global__ void score_decoded_kernel(const __grid_constant__ Geo g, const float* __restrict__ d0,
                                     const float* __restrict__ d1, const float* __restrict__ d2,
                                     float* __restrict__ score, float* __restrict__ prob, int32_t* __restrict__ cls) {
    const int lane = threadIdx.x & 31;
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wid >= g.rpi) return;
    const int off1 = g.head_rec[0], off2 = off1 + g.head_rec[1];
    const int h = wid >= off2 ? 2 : (wid >= off1 ? 1 : 0);
    const int rih = wid - (h == 0 ? 0 : (h == 1 ? off1 : off2));
    const float* r = (h == 0 ? d0 : (h == 1 ? d1 : d2)) + (size_t)rih * g.L;
    float best = 1.0f;
    int arg = 0;
    if (g.C > 0) warp_max_argmax(r + 5, g.C, lane, best, arg);
    if (lane == 0) {
        const float conf = r[4];
        score[wid] = g.C > 0 ? __fmul_rn(best, conf) : conf;
        prob[wid] = best;
        cls[wid] = arg;
    }
}

// ---- order-preserving compaction ------------------------------------------------------------------------
constexpr int kCompactThreads = 1024;

struct CompactSrc {
    const float* predicts;          // packed logits (mode 0)
    const float* decoded[3];        // decoded heads of one image (mode 1)
    const float* boxes[3];
    int mode;
};

__global__ void __launch_bounds__(kCompactThreads) compact_kernel(const __grid_constant__ Geo g, const CompactSrc s,
                                                                  float thresh, const float* __restrict__ score,
                                                                  const float* __restrict__ prob,
                                                                  const int32_t* __restrict__ cls, float* __restrict__ cand,
                                                                  int capacity, int32_t* __restrict__ counts,
                                                                  int32_t* status) {
    __shared__ int warp_tot[kCompactThreads / 32];
    __shared__ int head_total;
    const int n = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* sc = score + (size_t)n * g.rpi;
    const float* pr = prob + (size_t)n * g.rpi;
    const int32_t* cl = cls + (size_t)n * g.rpi;
    float* out = cand + (size_t)n * capacity * YV3_CAND_COLS;
    int base = 0;            // rows already written for this image (previous heads)
    int head_off = 0;
    for (int h = 0; h < 3; ++h) {
        const int cnt = g.head_rec[h];
        const int ipt = (cnt + kCompactThreads - 1) / kCompactThreads;
        const int lo = min(tid * ipt, cnt), hi = min(lo + ipt, cnt);
        int mine = 0;
        for (int i = lo; i < hi; ++i) mine += (sc[head_off + i] > thresh) ? 1 : 0;     // :62 strict, fp32
        // block exclusive scan of `mine`
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = warp_tot[lane];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += v;
            }
            warp_tot[lane] = wi - w;                  // exclusive prefix of warp totals
            if (lane == 31) head_total = wi;
        }
        __syncthreads();
        int pos = base + warp_tot[warp] + incl - mine;
        for (int i = lo; i < hi; ++i) {
            const float sv = sc[head_off + i];
            if (!(sv > thresh)) continue;
            if (pos < capacity) {
                float x0, y0, x1, y1, conf;
                if (s.mode == 0) {
                    const int b = i % g.B[h];
                    const int cell = i / g.B[h];
                    RecLoc p;
                    p.h = h; p.b = b; p.y = cell / g.W[h]; p.x = cell - p.y * g.W[h];
                    const float* r = s.predicts + ((size_t)n * g.rpi + record_of(g, h, p.y, p.x, b)) * g.L;
                    const BoxDec d = decode_box(g, p, __ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3), __ldg(r + 4));
                    x0 = d.x0; y0 = d.y0; x1 = d.x1; y1 = d.y1; conf = d.conf;
                } else {
                    const float4 bx = reinterpret_cast<const float4*>(s.boxes[h])[i];
                    x0 = bx.x; y0 = bx.y; x1 = bx.z; y1 = bx.w;
                    conf = s.decoded[h][(size_t)i * g.L + 4];
                }
                float4* o = reinterpret_cast<float4*>(out + (size_t)pos * YV3_CAND_COLS);
                o[0] = make_float4(__fdiv_rn(x0, g.fW[h]), __fdiv_rn(y0, g.fH[h]), __fdiv_rn(x1, g.fW[h]),
                                   __fdiv_rn(y1, g.fH[h]));                          // :65-68
                o[1] = make_float4(conf, pr[head_off + i], (float)cl[head_off + i], sv);
            }
            ++pos;
        }
        const int tot = head_total;
        if (tid == 0) counts[n * 3 + h] = tot;
        base += tot;
        head_off += cnt;
        __syncthreads();                              // warp_tot / head_total reused by the next head
    }
    if (tid == 0 && base > capacity) atomicOr((unsigned*)status, kStatusCapacity);
}

// ---- sort: (class asc, score desc, row asc) -----------------------------------------------------------
constexpr int kSortThreads = 1024;
constexpr int kSortSmemKeys = 16384;
constexpr unsigned long long kPadKey = ~0ULL;

__device__ __forceinline__ unsigned long long make_key(float cls, float score, int row) {
    // class in the top 14 bits, then the score so that LARGER scores sort first, then the row (18 bits):
    // equal scores keep their pre-sort order like Python's stable sorted() (:115)
    unsigned u = __float_as_uint(score);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);      // ascending-order image of the float
    u = ~u;                                              // descending
    const unsigned long long c = (unsigned long long)(unsigned)(int)cls & 0x3fffULL;
    return (c << 50) | ((unsigned long long)u << 18) | (unsigned long long)(row & 0x3ffff);
}

__device__ __forceinline__ void bitonic_sort(unsigned long long* keys, int n2, int tid) {
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < (n2 >> 1); i += kSortThreads) {
                const int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));
                const int hi = lo | j;
                const unsigned long long a = keys[lo], b = keys[hi];
                const bool up = (lo & k) == 0;
                if ((a > b) == up) { keys[lo] = b; keys[hi] = a; }
            }
            __syncthreads();
        }
    }
}

__global__ void __launch_bounds__(kSortThreads) sort_kernel(const float* __restrict__ cand, const int32_t* __restrict__ count,
                                                            int count_cols, int capacity, int cap2,
                                                            unsigned long long* __restrict__ gkeys, float4* __restrict__ sbox,
                                                            int32_t* __restrict__ sidx) {
    extern __shared__ __align__(16) unsigned long long skeys[];
    const int n = blockIdx.x, tid = threadIdx.x;
    int K = 0;
    for (int c = 0; c < count_cols; ++c) K += count[n * count_cols + c];
    K = max(0, min(K, capacity));
    int n2 = 1;
    while (n2 < K) n2 <<= 1;
    if (n2 < 2) n2 = 2;
    const float* rows = cand + (size_t)n * capacity * YV3_CAND_COLS;
    unsigned long long* gk = gkeys + (size_t)n * cap2;
    unsigned long long* keys = (n2 <= kSortSmemKeys) ? skeys : gk;
    for (int i = tid; i < n2; i += kSortThreads) {
        unsigned long long key = kPadKey;
        if (i < K) {
            const float4 t = reinterpret_cast<const float4*>(rows + (size_t)i * YV3_CAND_COLS)[1];
            key = make_key(t.z, t.w, i);
        }
        keys[i] = key;
    }
    __syncthreads();
    bitonic_sort(keys, n2, tid);
    for (int i = tid; i < K; i += kSortThreads) {
        const unsigned long long key = keys[i];
        const int row = (int)(key & 0x3ffffULL);
        if (keys != gk) gk[i] = key;
        sidx[(size_t)n * capacity + i] = row;
        sbox[(size_t)n * capacity + i] = reinterpret_cast<const float4*>(rows + (size_t)row * YV3_CAND_COLS)[0];
    }
}

// ---- greedy NMS per (class, image) segment ---------------------------------------------------------------
constexpr int kNmsThreads = 256;
constexpr int kNmsChunk = 256;
constexpr int kKeptSmem = 2048;

struct BoxD {
    double x0, y0, x1, y1, area;
};
__device__ __forceinline__ BoxD to_boxd(const float4 b) {
    BoxD o;
    o.x0 = (double)b.x; o.y0 = (double)b.y; o.x1 = (double)b.z; o.y1 = (double)b.w;
    o.area = __dmul_rn(__dsub_rn(o.x1, o.x0), __dsub_rn(o.y1, o.y0));     // (x1 - x0) * (y1 - y0)
    return o;
}
// _cal_iou(box, truth) > thresh  (:132-145, :124): w/h early-out, union = (area_a + area_b) - inter
__device__ __forceinline__ bool suppresses(const BoxD& a, const BoxD& b, double thresh) {
    const double w = __dsub_rn(fmin(a.x1, b.x1), fmax(a.x0, b.x0));
    const double h = __dsub_rn(fmin(a.y1, b.y1), fmax(a.y0, b.y0));
    if (w <= 0.0 || h <= 0.0) return 0.0 > thresh;
    const double inter = __dmul_rn(w, h);
    const double uni = __dsub_rn(__dadd_rn(a.area, b.area), inter);
    return __ddiv_rn(inter, uni) > thresh;
}

__global__ void __launch_bounds__(kNmsThreads) nms_kernel(const unsigned long long* __restrict__ gkeys,
                                                          const int32_t* __restrict__ count, int count_cols,
                                                          int capacity, int cap2, double thresh,
                                                          const float4* __restrict__ sbox, const int32_t* __restrict__ sidx,
                                                          float4* __restrict__ kspill, unsigned char* __restrict__ keep) {
    __shared__ float4 kept_s[kKeptSmem];
    __shared__ float4 chunk_s[kNmsChunk];
    __shared__ unsigned long long mask_s[kNmsChunk][kNmsChunk / 64];
    __shared__ unsigned char alive_s[kNmsChunk];
    __shared__ unsigned char kept_flag_s[kNmsChunk];
    __shared__ int seg_s[2];
    __shared__ int nkept_s;
    const int c = blockIdx.x, n = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int K = 0;
    for (int q = 0; q < count_cols; ++q) K += count[n * count_cols + q];
    K = max(0, min(K, capacity));
    if (K == 0) return;
    const unsigned long long* keys = gkeys + (size_t)n * cap2;
    if (tid < 2) {
        // first sorted position whose class field is >= c + tid
        const unsigned long long want = (unsigned long long)(c + tid);
        int lo = 0, hi = K;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((keys[mid] >> 50) < want) lo = mid + 1; else hi = mid;
        }
        seg_s[tid] = lo;
    }
    if (tid == 0) nkept_s = 0;
    __syncthreads();
    const int s0 = seg_s[0], s1 = seg_s[1];
    if (s0 >= s1) return;
    const float4* boxes = sbox + (size_t)n * capacity;
    float4* spill = kspill + (size_t)n * capacity + s0;
    const int32_t* rowof = sidx + (size_t)n * capacity;
    unsigned char* keep_n = keep + (size_t)n * capacity;

    for (int c0 = s0; c0 < s1; c0 += kNmsChunk) {
        const int cnt = min(kNmsChunk, s1 - c0);
        const int nk = nkept_s;
        // phase 1: this thread's candidate against every box kept so far
        bool alive = tid < cnt;
        float4 mine_f = make_float4(0.f, 0.f, 0.f, 0.f);
        BoxD mine = {};
        if (alive) {
            mine_f = boxes[c0 + tid];
            mine = to_boxd(mine_f);
            chunk_s[tid] = mine_f;
        }
        for (int q = 0; q < nk; ++q) {
            if (!__any_sync(0xffffffffu, alive)) break;
            const float4 kb = (q < kKeptSmem) ? kept_s[q] : spill[q];
            if (alive && suppresses(to_boxd(kb), mine, thresh)) alive = false;
        }
        alive_s[tid] = alive ? 1 : 0;
        __syncthreads();
        // phase 2: bit matrix inside the chunk: row i, bit j (j > i) <=> i suppresses j
        {
            unsigned long long w0 = 0, w1 = 0, w2 = 0, w3 = 0;
            if (alive) {
                for (int j = tid + 1; j < cnt; ++j) {
                    if (!alive_s[j]) continue;
                    if (suppresses(mine, to_boxd(chunk_s[j]), thresh)) {
                        const unsigned long long bit = 1ULL << (j & 63);
                        const int w = j >> 6;
                        if (w == 0) w0 |= bit; else if (w == 1) w1 |= bit; else if (w == 2) w2 |= bit; else w3 |= bit;
                    }
                }
            }
            mask_s[tid][0] = w0; mask_s[tid][1] = w1; mask_s[tid][2] = w2; mask_s[tid][3] = w3;
        }
        __syncthreads();
        // sweep: warp 0 walks the chunk in order; lane w (< 4) carries word w of the removed set
        if (warp == 0) {
            unsigned long long removed = 0;
            for (int i = 0; i < cnt; ++i) {
                const unsigned long long word = __shfl_sync(0xffffffffu, removed, i >> 6);
                const bool k = alive_s[i] && !((word >> (i & 63)) & 1ULL);
                if (k && lane < 4) removed |= mask_s[i][lane];
                if (lane == 0) kept_flag_s[i] = k ? 1 : 0;
            }
        }
        __syncthreads();
        // append the chunk's survivors to the kept list (order kept) and publish the flags
        if (warp == 0) {
            int total = nk;
            for (int b = 0; b < cnt; b += 32) {
                const int i = b + lane;
                const bool k = i < cnt && kept_flag_s[i];
                const unsigned m = __ballot_sync(0xffffffffu, k);
                if (k) {
                    const int pos = total + __popc(m & ((1u << lane) - 1u));
                    if (pos < kKeptSmem) kept_s[pos] = chunk_s[i]; else spill[pos] = chunk_s[i];
                }
                total += __popc(m);
            }
            if (lane == 0) nkept_s = total;
        }
        if (tid < cnt) keep_n[rowof[c0 + tid]] = kept_flag_s[tid];
        __syncthreads();
    }
}

// ---- launchers ------------------------------------------------------------------------------------------
static TileLaunch score_tiles(const Geo& g, const float* predicts, int n) {
    TileLaunch q;
    q.total_rec = (long long)n * g.rpi;
    long long tr = (long long)(45056 / ((size_t)g.L * 4));       // ~44 KB per CTA: five CTAs per SM
    tr = min(tr, (long long)kScoreThreads * 4);
    if (tr > kScoreThreads) tr = tr / kScoreThreads * kScoreThreads;
    tr &= ~3LL;
    if (tr < 4) tr = 4;
    q.tr = (int)tr;
    q.num_tiles = (int)((q.total_rec + tr - 1) / tr);
    q.tma = (((uintptr_t)predicts & 15u) == 0) ? 1 : 0;
    return q;
}

cudaError_t launch_filter_packed(const Geo& g, const float* predicts, int n, float thresh, float* cand,
                                 int capacity, int32_t* counts, int32_t* status, float* score, float* prob,
                                 int32_t* cls, int sm_count, cudaStream_t stream) {
    (void)sm_count;
    const TileLaunch q = score_tiles(g, predicts, n);
    const size_t smem = (size_t)q.tr * g.L * 4 + 128;
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t err = cudaFuncSetAttribute(score_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    score_packed_kernel<<<q.num_tiles, kScoreThreads, smem, stream>>>(g, predicts, q, score, prob, cls);
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    CompactSrc s;
    memset(&s, 0, sizeof(s));
    s.predicts = predicts;
    s.mode = 0;
    compact_kernel<<<n, kCompactThreads, 0, stream>>>(g, s, thresh, score, prob, cls, cand, capacity, counts, status);
    return cudaGetLastError();
}

cudaError_t launch_filter_decoded(const Geo& g, const float* const decoded[3], const float* const boxes[3],
                                  float thresh, float* cand, int capacity, int32_t* counts, int32_t* status,
                                  float* score, float* prob, int32_t* cls, int sm_count, cudaStream_t stream) {
    (void)sm_count;
    const int threads = 256;
    const long long warps = g.rpi;
    const int blocks = (int)((warps * 32 + threads - 1) / threads);
    score_decoded_kernel<<<blocks, threads, 0, stream>>>(g, decoded[0], decoded[1], decoded[2], score, prob, cls);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    CompactSrc s;
    memset(&s, 0, sizeof(s));
    for (int h = 0; h < 3; ++h) { s.decoded[h] = decoded[h]; s.boxes[h] = boxes[h]; }
    s.mode = 1;
    compact_kernel<<<1, kCompactThreads, 0, stream>>>(g, s, thresh, score, prob, cls, cand, capacity, counts, status);
    return cudaGetLastError();
}

int nms_cap2(int capacity) {
    int c = 2;
    while (c < capacity) c <<= 1;
    return c;
}

cudaError_t launch_nms(const float* cand, const int32_t* count, int count_cols, int n, int capacity,
                       int num_classes, double thresh, unsigned char* keep, const PostWs& ws,
                       cudaStream_t stream) {
    const int cap2 = nms_cap2(capacity);
    cudaError_t err = cudaMemsetAsync(keep, 0, (size_t)n * capacity, stream);
    if (err != cudaSuccess) return err;
    const size_t smem = (size_t)min(cap2, kSortSmemKeys) * 8;
    err = cudaFuncSetAttribute(sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    sort_kernel<<<n, kSortThreads, smem, stream>>>(cand, count, count_cols, capacity, cap2, ws.keys, ws.sbox, ws.sidx);
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    dim3 grid((unsigned)max(num_classes, 1), (unsigned)n);
    nms_kernel<<<grid, kNmsThreads, 0, stream>>>(ws.keys, count, count_cols, capacity, cap2, thresh, ws.sbox, ws.sidx,
                                                 ws.kspill, keep);
    return cudaGetLastError();
}

}  // namespace yv3
