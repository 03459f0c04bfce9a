// Score thresholding and class-aware greedy NMS (yolov3/yolov3_post_process.py).
//
//   filter  _filter_single_head_boxes :44-76   score = max_c p_c * conf (fp32), class = FIRST arg-max,
//           keep iff score > fp32(thresh) (strict); rows [x0/W, y0/H, x1/W, y1/H, conf, prob, cls, score]
//           in head /8,/16,/32 order, raster (y, x, b) inside a head.
//   nms     _apply_nms :108-129, _cal_iou :132-145   stable sort by score descending, a later box is
//           dropped iff an earlier KEPT box of the same class has IoU_f64 > thresh (strict).
//
// Kernels:
//   score_tile_kernel     one sweep over the packed logits (HBM-bound, 4*A*L bytes/image): tile in shared
//                         memory by bulk copy, one thread per record, approximate early-out, exact scores for
//                         the few records near or above the threshold, survivors ranked inside the tile.
//   gather_kernel         one CTA per image: prefix over the tiles' counts, candidate rows written in the
//                         reference's order; survivors' boxes are decoded from 5 logits.
//   score_decoded_kernel / compact_kernel   the same for already decoded heads of one image (filter_boxes' inputs).
//   sort_kernel           one CTA per image: 64-bit keys (class | ~score | row) bitonic-sorted in shared
//                         memory (global memory beyond 16 Ki candidates).
//   nms_kernel<T>         persistent CTAs over the queued long segments, T-candidate chunks; each chunk is first
//                         tested against the boxes kept so far, then resolved internally with a
//                         shared-memory IoU bit matrix and a warp sweep.  IoU is fp64, uncontracted, in the
//                         reference's operation order.
#include <cooperative_groups.h>
#include <stdlib.h>
#include <string.h>

#include <stdio.h>

#include "sweep_parts.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kScoreThreads = 128;

// ---- fused score + threshold over the packed logits ---------------------------------------------------------
// One CTA per tile of <= 128 consecutive records (<= 2 images), one thread per record:
//  1. the tile lands in shared memory (bulk copy);
//  2. every thread bounds its record's score from above: score = fl(clip(max_c p_c) * conf) <= conf, and
//     max_c p_c = 1 / sum_k exp(z_k - max z) where the sum is first taken with ex2.approx (4 instructions per
//     class instead of ~16).  A record whose bound, widened by 1e-3, stays below the threshold cannot pass
//     (:62 is strict) and is dropped; the others ("maybe", a few per tile) are queued;
//  3. the queued records are scored exactly -- the same class_stats / class_best as the decode kernel, one record per
//     thread of the FIRST warps, so the expensive path runs on dense lanes;
//  4. survivors are ranked inside the tile per (image slot, head) with warp ballots and written to the tile's
//     staging rows (score, prob, class, record-in-head); six counts per tile.
// gather_kernel then turns counts into the reference's order (head /8,/16,/32, raster inside a head) per image and
// writes the candidate rows, decoding each survivor's box from its five logits.

// A tile is the <= 32 consecutive records of ONE WARP: the warp lands it in its own slice of shared memory with
// its own bulk copy and mbarrier, scores, ranks and stages it without ever meeting the other warps of the CTA (the
// CTA only bundles warps so that their buffers fill the SM's shared memory).
__global__ void __launch_bounds__(kScoreThreads) score_tile_kernel(const __grid_constant__ Geo g,
                                                                   const float* __restrict__ predicts, const ScoreTiles q,
                                                                   float thresh, StageRow* __restrict__ stage,
                                                                   int32_t* __restrict__ tcount) {
    extern __shared__ __align__(128) unsigned char ssm[];
    __shared__ uint64_t bar_s[kScoreThreads / 32];
    const int L = g.L, tr = q.tr;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int unit = blockIdx.x * q.wpc + warp;
    if (warp >= q.wpc || unit >= q.num_tiles) return;
    float* tile = (float*)ssm + (size_t)warp * q.tile_floats;
    const long long rec0 = (long long)unit * tr;
    const int nrec = (int)min((long long)tr, q.total_rec - rec0);
    {
        const float* src = predicts + (size_t)rec0 * L;
        const int nfl = nrec * L;
        const unsigned bulk = q.tma ? (unsigned)(((size_t)nfl * 4) & ~(size_t)15) : 0u;
        if (bulk) {
            uint64_t* bar = &bar_s[warp];
            if (lane == 0) {
                mbar_init(bar, 1);
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                mbar_expect_tx(bar, bulk);
                bulk_g2s(tile, src, bulk, bar);
            }
            for (int i = (int)(bulk / 4) + lane; i < nfl; i += 32) tile[i] = __ldg(src + i);
            __syncwarp();                              // barrier init visible to every lane
            mbar_wait(bar, 0);
        } else {
            for (int i = lane; i < nfl; i += 32) tile[i] = __ldg(src + i);
        }
        __syncwarp();
    }

    // the lane's record: image slot of the tile (a tile touches at most two images), head / cell / anchor, confidence
    const int n_first = (int)(rec0 / g.rpi);
    int r_img = (int)(rec0 - (long long)n_first * g.rpi) + lane, slot = 0;
    if (r_img >= g.rpi) { r_img -= g.rpi; slot = 1; }
    RecLoc p = {0, 0, 0, 0};
    float conf = 0.f;
    if (lane < nrec) {
        p = locate(g, r_img);
        conf = clip_ref(sigmoid_ref(tile[(size_t)lane * L + 4]), g.lo, g.hi);
    }
    score_tile_records(g, q, tile, unit, lane, nrec, slot, p, conf, thresh, stage, tcount);
}

// One CTA per image: per head, the counts of the tiles that hold the image's records of that head go to shared memory
// and an exclusive prefix puts them in the reference's order (head /8,/16,/32, raster inside a head: tiles in order,
// staged order inside a tile); then one THREAD per candidate row finds its tile by binary search and copies the row the
// sweep finished (:65-75; sweep_parts.cuh) -- the tensor is not touched again.  `src` says where head h's records and tiles live: the packed tensor
// (all three heads share one tile range) or one tensor per head.
struct GatherSrc {
    const float* pred[3];
    int rpi[3], rpr[3], seg[3];     // records per image / per 32-row / first record of the head in a row, in that tensor
    int tbase[3];                   // first tile of that tensor in stage / tcount
};
constexpr int kGatherThreads = 512;
__global__ void __launch_bounds__(kGatherThreads) gather_kernel(const __grid_constant__ Geo g, const __grid_constant__ GatherSrc src,
                                                                int tr, const StageRow* __restrict__ stage,
                                                                const int32_t* __restrict__ tcount, float* __restrict__ cand,
                                                                int capacity, int32_t* __restrict__ counts, int32_t* status) {
    __shared__ int wsum_s[kGatherThreads / 32];
    __shared__ int carry_s;
    __shared__ int head_base_s[4];
    __shared__ int t0_s[3], nt_s[3], off_s[3];
    extern __shared__ int gsm[];                   // per head list: cnt_s[nt], before_s[nt], pre_s[nt] (concatenated)
    const int n = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        int off = 0;
        for (int h = 0; h < 3; ++h) {
            const long long img0 = (long long)n * src.rpi[h];
            const int t0 = (int)(img0 / tr), t1 = (int)((img0 + src.rpi[h] - 1) / tr);
            t0_s[h] = t0; nt_s[h] = t1 - t0 + 1; off_s[h] = off;
            off += 3 * (t1 - t0 + 1);
        }
    }
    __syncthreads();
    for (int h = 0; h < 3; ++h) {
        const int nt = nt_s[h];
        int* cnt_s = gsm + off_s[h];
        int* before_s = cnt_s + nt;
        for (int ti = tid; ti < nt; ti += kGatherThreads) {
            const int tl = t0_s[h] + ti;                                               // tile inside its tensor
            const int sl = n - (int)(((long long)tl * tr) / src.rpi[h]);              // 0 or 1: this image's slot in the tile
            const int32_t* tc = tcount + (size_t)(src.tbase[h] + tl) * 8;
            int before = 0;
            for (int gi = 0; gi < sl * 3 + h; ++gi) before += tc[gi];                  // staged rows ahead of this group
            cnt_s[ti] = tc[sl * 3 + h];
            before_s[ti] = before;
        }
    }
    __syncthreads();
    int total_all = 0;
    for (int h = 0; h < 3; ++h) {
        const int nt = nt_s[h];
        const int* cnt_s = gsm + off_s[h];
        int* pre_s = gsm + off_s[h] + 2 * nt;
        if (tid == 0) carry_s = 0;
        __syncthreads();
        for (int b = 0; b < nt; b += kGatherThreads) {
            const int ti = b + tid;
            const int c = (ti < nt) ? cnt_s[ti] : 0;
            int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) wsum_s[warp] = incl;
            __syncthreads();
            int wbase = carry_s;
            for (int w = 0; w < warp; ++w) wbase += wsum_s[w];
            if (ti < nt) pre_s[ti] = wbase + incl - c;
            __syncthreads();
            if (tid == kGatherThreads - 1) carry_s = wbase + incl;
            __syncthreads();
        }
        if (tid == 0) { head_base_s[h] = total_all; counts[n * 3 + h] = carry_s; }
        total_all += carry_s;
        __syncthreads();
    }
    if (tid == 0) {
        head_base_s[3] = total_all;
        if (total_all > capacity) atomicOr((unsigned*)status, kStatusCapacity);
    }
    __syncthreads();
    float* out = cand + (size_t)n * capacity * YV3_CAND_COLS;
    const int nrows = min(total_all, capacity);
    for (int pos = tid; pos < nrows; pos += kGatherThreads) {
        const int h = (pos >= head_base_s[2]) ? 2 : ((pos >= head_base_s[1]) ? 1 : 0);
        const int rel = pos - head_base_s[h];
        const int nt = nt_s[h];
        const int* cnt_s = gsm + off_s[h];
        const int* before_s = cnt_s + nt;
        const int* pre = cnt_s + 2 * nt;
        // last tile whose exclusive prefix is <= rel and that holds rows (prefix strictly increases past it)
        int lo = 0, hi = nt - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (pre[mid] <= rel) lo = mid; else hi = mid - 1;
        }
        // tiles with zero rows share the prefix of their successor: step back to the one that contains rel
        while (rel - pre[lo] >= cnt_s[lo]) --lo;
        const int e = rel - pre[lo];
        const StageRow sv = stage[(size_t)(src.tbase[h] + t0_s[h] + lo) * tr + before_s[lo] + e];
        float4* o = reinterpret_cast<float4*>(out + (size_t)pos * YV3_CAND_COLS);
        o[0] = sv.box;
        o[1] = sv.tail;
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

// ---- fp64 IoU test of the NMS (shared by the fused small-segment pass and the chunked kernel) -------------
struct BoxD {
    double x0, y0, x1, y1, area;
};
__device__ __forceinline__ BoxD to_boxd(const float4 b) {
    BoxD o;
    o.x0 = (double)b.x; o.y0 = (double)b.y; o.x1 = (double)b.z; o.y1 = (double)b.w;
    o.area = __dmul_rn(__dsub_rn(o.x1, o.x0), __dsub_rn(o.y1, o.y0));     // (x1 - x0) * (y1 - y0)
    return o;
}
// _cal_iou(box, truth) > thresh  (:132-145, :124) for two boxes that are known to overlap: union = (area_a + area_b) -
// inter; fl(inter / uni) > thresh is decided without the division unless the quotient lies within 2^-48 (relative)
// of the threshold (the products are off by at most 2^-52 each), then by the division itself.
// "Known to overlap": overlaps_f32, defined below, is the complement of the reference's w <= 0 or h <= 0 early-out
// (:139-142).  The decision is taken from the fp32 rows: the coordinates are fp32 values, so min / max
// can be taken in fp32 and converted -- (double)fminf(a, b) == fmin((double)a, (double)b) -- which replaces four
// double min/max (a compare and two selects each) by conversions of results the overlap test has already computed.
__device__ __forceinline__ double box_area_d(const float4& b) {
    return __dmul_rn(__dsub_rn((double)b.z, (double)b.x), __dsub_rn((double)b.w, (double)b.y));    // (x1 - x0) * (y1 - y0)
}
__device__ __forceinline__ bool suppresses_overlapping(const float4& a, double area_a, const float4& b, double area_b,
                                                       double thresh) {
    const double w = __dsub_rn((double)fminf(a.z, b.z), (double)fmaxf(a.x, b.x));
    const double h = __dsub_rn((double)fminf(a.w, b.w), (double)fmaxf(a.y, b.y));
    const double inter = __dmul_rn(w, h);
    const double uni = __dsub_rn(__dadd_rn(area_a, area_b), inter);
    const double tu = __dmul_rn(thresh, uni);
    if (inter > __dmul_rn(tu, 1.0000000000000036)) return true;          // 1 + 2^-48
    if (inter < __dmul_rn(tu, 0.9999999999999964)) return false;         // 1 - 2^-48
    return __ddiv_rn(inter, uni) > thresh;
}

// The same decision, cheaper: almost every overlapping pair is far from the threshold, and for those fp32 arithmetic
// on the fp32 coordinates already decides.  inter32 and uni32 below carry a relative error of at most 3 * 2^-24 and
// 12 * 2^-24 (differences and products of exact fp32 inputs, each rounded once; uni >= (a + b) / 2 >= inter bounds the
// cancellation in (a + b) - inter), the threshold product 2 more: under 2^-19 in all against the exact real quotient,
// and the reference's own fp64 roundings are 2^-52.  So outside a band of +-2^-18 (relative) around the threshold the
// fp32 comparison IS the reference's decision; inside it -- or when a product left the normal fp32 range -- the exact
// fp64 expression decides.  thr_lo / thr_hi: float(thresh * (1 -+ 2^-18)) rounded outward (NmsThresh below).
struct NmsThresh {
    double t;
    float lo, hi;
};
__host__ __device__ inline NmsThresh make_nms_thresh(double t) {
    NmsThresh q;
    q.t = t;
    const double band = 3.814697265625e-06;                              // 2^-18
    const double a = t * (1.0 - band), b = t * (1.0 + band);
    const double dlo = a < b ? a : b, dhi = a < b ? b : a;               // a negative threshold swaps them
    float lo = (float)dlo, hi = (float)dhi;
    if ((double)lo > dlo) lo = nextafterf(lo, -INFINITY);
    if ((double)hi < dhi) hi = nextafterf(hi, INFINITY);
    q.lo = lo;
    q.hi = hi;
    return q;
}
__device__ __forceinline__ bool suppresses_overlapping_fast(const float4& a, const float4& b, const NmsThresh& q) {
    const float w = __fsub_rn(fminf(a.z, b.z), fmaxf(a.x, b.x));
    const float h = __fsub_rn(fminf(a.w, b.w), fmaxf(a.y, b.y));
    const float inter = __fmul_rn(w, h);
    const float aa = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
    const float ab = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
    const float uni = __fsub_rn(__fadd_rn(aa, ab), inter);
    const bool normal = inter >= 1e-30f && aa >= 1e-30f && ab >= 1e-30f && uni <= 1e30f;     // false for NaN as well
    if (normal && inter > __fmul_rn(q.hi, uni) && q.hi >= 0.f) return true;
    if (normal && inter < __fmul_rn(q.lo, uni) && q.lo >= 0.f) return false;
    return suppresses_overlapping(a, box_area_d(a), b, box_area_d(b), q.t);
}

// ---- sort: (class asc, score desc, row asc) -----------------------------------------------------------
// 1024 threads: the gather -> sort -> NMS chain of a step must stay well below a sweep's duration or the post-process
// stream becomes the step's bottleneck (measured at cfg2: 256 threads 103 us/step like 1024, 128 threads 126 us/step)
constexpr int kSortThreads = 1024;
constexpr int kSortInlineWarps = 8;     // warps of the last sort CTA that serve a short queue of medium segments themselves
constexpr int kSortInlineJobs = 64;     // ... up to this many segments
constexpr int kSortSmemKeys = 4096;      // lists up to this size are sorted in shared memory (45 KB per CTA)
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

// Lists of at most kSortThreads keys: one key per thread in a register; compare-exchange partners closer than a warp
// are reached by shuffles, the others through shared memory (15 of the 55 steps at 1024 keys).  `keys` must be shared
// memory; on return keys[0..n2) is sorted.
__device__ __forceinline__ void bitonic_sort_regs(unsigned long long* keys, int n2, int tid) {
    unsigned long long mine = (tid < n2) ? keys[tid] : kPadKey;
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            unsigned long long other;
            if (j >= 32) {
                __syncthreads();                      // partners have read the previous round's values
                if (tid < n2) keys[tid] = mine;
                __syncthreads();
                other = (tid < n2) ? keys[tid ^ j] : kPadKey;
            } else {
                other = __shfl_xor_sync(0xffffffffu, mine, j);
            }
            const bool lower = (tid & j) == 0, up = (tid & k) == 0;
            const bool take_min = lower == up;
            const unsigned long long mn = mine < other ? mine : other, mx = mine < other ? other : mine;
            mine = take_min ? mn : mx;
        }
    }
    __syncthreads();
    if (tid < n2) keys[tid] = mine;
    __syncthreads();
}

// The branch-free fp32 pair test of the NMS kernels (warp_nms_tile and its relatives) is trusted when every area of the segment is at least kWarpNmsAreaLo (and at most 1e29) and
// the threshold at least kWarpNmsThreshLo: an intersection that leaves the normal fp32 range (below 1e-30, where the
// relative error bound of the product no longer holds) is then still 6 orders of magnitude below lo * union >= 1e-24,
// i.e. on the "no" side whatever its rounding; disjoint boxes (inter <= 0) are on the "no" side because lo * union > 0.
constexpr float kWarpNmsAreaLo = 1e-18f;
constexpr float kWarpNmsThreshLo = 1e-6f;
constexpr int kNmsMedium = 256;         // threads = chunk size of the NMS kernel for segments of 33..256 candidates
constexpr int kNmsLarge = 1024;         // ... and for longer ones
constexpr int kFuseMaxSeg = 32;         // class segments up to one warp wide are resolved inside the sort kernel
constexpr int kBucketMaxK = 65535;      // bucket path of the sort kernel: lists up to this size (arrival order is 16 bits) ...
constexpr int kBucketClasses = 1024;    // ... with class ids below this (thread t owns classes 4t .. 4t + 3) and <= 256 per class
constexpr int kClsPerThread = kBucketClasses / kSortThreads;
constexpr int kWarpNmsSlots = 8;        // candidates per lane of the warp-per-segment NMS
constexpr int kWarpNmsMax = 32 * kWarpNmsSlots;    // = kNmsMedium: the longest segment one warp resolves

// Lists of 1025..n2s/2 keys in shared memory (one long list per image: the heads with a handful of classes): a bin sort
// instead of the bitonic network (91 rounds over 8192 keys: 100 us per step at the single-class stress shape).  The upper
// 46 bits of a key (class, inverted score) are mapped linearly from [min, max] onto 1024 bins -- monotone, so the bins are
// in key order -- counted, prefixed and scattered into the upper half of the key buffer; every key then ranks itself
// among the keys of its own bin (~8 of them) and is written to its final place in the lower half.  A distribution that
// piles more than 256 keys into one bin (all scores equal ...) is left to the bitonic sort: returns false, keys untouched.
__device__ __forceinline__ bool bin_sort_keys(unsigned long long* keys, const int K, int* hist_s, unsigned short* pos_s,
                                              const int tid) {
    static_assert(kBucketClasses == kSortThreads, "one bin per thread");
    __shared__ unsigned long long red_s[2][kSortThreads / 32];
    __shared__ int wtot2_s[kSortThreads / 32];
    const int lane = tid & 31, warp = tid >> 5;
    unsigned long long mn = ~0ULL, mx = 0ULL;
    for (int i = tid; i < K; i += kSortThreads) {
        const unsigned long long u = keys[i] >> 18;
        mn = u < mn ? u : mn;
        mx = u > mx ? u : mx;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
        mn = a < mn ? a : mn;
        mx = b > mx ? b : mx;
    }
    if (lane == 0) { red_s[0][warp] = mn; red_s[1][warp] = mx; }
    hist_s[tid] = 0;                                              // kBucketClasses == kSortThreads bins
    __syncthreads();
    mn = red_s[0][lane];
    mx = red_s[1][lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long a = __shfl_xor_sync(0xffffffffu, mn, o), b = __shfl_xor_sync(0xffffffffu, mx, o);
        mn = a < mn ? a : mn;
        mx = b > mx ? b : mx;
    }
    const double scale = mx > mn ? (double)(kBucketClasses - 1) / (double)(mx - mn) : 0.0;
    auto bin_of = [&](unsigned long long key) { return (int)((double)((key >> 18) - mn) * scale); };
    for (int i = tid; i < K; i += kSortThreads) pos_s[i] = (unsigned short)min(atomicAdd(&hist_s[bin_of(keys[i])], 1), 65535);
    __syncthreads();
    const int cnt = hist_s[tid];
    if (__syncthreads_or(cnt > kWarpNmsMax ? 1 : 0)) return false;
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) wtot2_s[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < warp; ++w) base += wtot2_s[w];
    hist_s[tid] = base + incl - cnt;                              // bin -> first position
    __syncthreads();
    unsigned long long* up = keys + K;
    for (int i = tid; i < K; i += kSortThreads) {
        const unsigned long long key = keys[i];
        up[hist_s[bin_of(key)] + pos_s[i]] = key;
    }
    __syncthreads();
    for (int p = tid; p < K; p += kSortThreads) {
        const unsigned long long key = up[p];
        const int b = bin_of(key);
        const int o0 = hist_s[b], o1 = (b + 1 < kBucketClasses) ? hist_s[b + 1] : K;
        int rank = 0;
        for (int q = o0; q < o1; ++q) rank += (up[q] < key) ? 1 : 0;
        keys[o0 + rank] = key;                                    // keys are unique (the row is part of the key)
    }
    __syncthreads();
    return true;
}

// One CTA per image: sort keys (class asc, score desc, row asc) in shared memory, publish the sorted order, then --
// for lists of at most kFuseMaxK candidates -- resolve every class segment of at most 32 candidates right here, one
// warp per segment: lane j holds candidate j of the segment, the kept candidates are visited in order, each is
// broadcast from shared memory and tested against the later lanes in fp64 (_apply_nms :117-128).  Longer segments
// are left to nms_kernel.
__device__ __forceinline__ void sort_body(const float* __restrict__ cand, const int32_t* __restrict__ count,
                                          int count_cols, int capacity, int cap2, int n2s, double thresh,
                                          unsigned long long* __restrict__ gkeys, float4* __restrict__ sbox,
                                          int32_t* __restrict__ sidx, unsigned char* __restrict__ keep,
                                          int2* __restrict__ wl, int2* __restrict__ wl_large,
                                          int32_t* __restrict__ wl_count, const bool many_classes) {
    extern __shared__ __align__(16) unsigned long long skeys[];
    const NmsThresh nq = make_nms_thresh(thresh);
    const int n = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int K = 0;
    for (int c = 0; c < count_cols; ++c) K += count[n * count_cols + c];
    K = max(0, min(K, capacity));
    int n2 = 1;
    while (n2 < K) n2 <<= 1;
    if (n2 < 2) n2 = 2;
    const float* rows = cand + (size_t)n * capacity * YV3_CAND_COLS;
    unsigned long long* gk = gkeys + (size_t)n * cap2;
    unsigned long long* keys = (n2 <= n2s) ? skeys : gk;           // short lists live in shared memory, long ones in gk
    __syncthreads();
    // ---- bucket path: no class of this image has more than kWarpNmsMax (256) candidates -- ~6 per class at cfg2, ~133
    // at the dense stress workload (10 647 candidates per image over 80 classes).  Sorting by (class, score desc, row) is
    // then a counting sort by class -- a shared-memory histogram, a prefix over the classes, a scatter -- followed by a
    // rank sort inside each bucket by one warp (up to 8 keys per lane).  A bucket of at most 32 goes straight on to the
    // greedy NMS with the candidates in the warp's lanes; a longer one is published (sorted keys, rows, boxes) and queued
    // for nms_warp_kernel.  No 16 384-key bitonic sort (380 us per step at the stress workload).
    int* hist_s = reinterpret_cast<int*>(skeys + n2s);                                // [kBucketClasses] counts -> offsets
    // (with a handful of classes a long list cannot have <= 256 per class: no attempt, the histogram would be 8000
    // atomics on one word)
    if (K <= kBucketMaxK && (many_classes || K <= 8 * kWarpNmsMax)) {
        // The shared-memory reservation is sized for SHORT lists (n2s <= 4096 keys, 45 KB: a sort CTA finds room on an SM
        // next to the sweep's CTAs; with 16 384 keys it needed 164 KB and pushed ~15 of them out for its whole life).
        // A longer list keeps its bucketed keys in gk and the arrival order in the (still unused) sorted-row table.
        unsigned long long* bkey = keys;                                              // [n2] keys, bucketed
        unsigned short* pos_s = (n2 <= n2s) ? reinterpret_cast<unsigned short*>(hist_s + kBucketClasses)
                                            : reinterpret_cast<unsigned short*>(sidx + (size_t)n * capacity);
        __shared__ int wtot_s[kSortThreads / 32];
        __shared__ int maxc_s;                                                        // highest class that has candidates
        for (int c = tid; c < kBucketClasses; c += kSortThreads) hist_s[c] = 0;
        if (tid == 0) maxc_s = -1;
        __syncthreads();
        bool bad = false;
        for (int i = tid; i < K; i += kSortThreads) {
            const float4 t = reinterpret_cast<const float4*>(rows + (size_t)i * YV3_CAND_COLS)[1];
            const int c = (int)(make_key(t.z, t.w, i) >> 50);
            if (c >= kBucketClasses) bad = true;
            else pos_s[i] = (unsigned short)min(atomicAdd(&hist_s[c], 1), 65535);
        }
        __syncthreads();
        int cnt[kClsPerThread], mycount = 0;
        bool over = false;
#pragma unroll
        for (int q = 0; q < kClsPerThread; ++q) {
            cnt[q] = hist_s[tid * kClsPerThread + q];
            mycount += cnt[q];
            over = over || cnt[q] > kWarpNmsMax;
        }
        if (mycount > 0) atomicMax(&maxc_s, tid * kClsPerThread + kClsPerThread - 1);
        if (!__syncthreads_or((bad || over) ? 1 : 0)) {
            // exclusive prefix over the classes: thread t owns classes 4t .. 4t + 3
            int incl = mycount;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            if (lane == 31) wtot_s[warp] = incl;
            __syncthreads();
            int base = 0;
            for (int w = 0; w < warp; ++w) base += wtot_s[w];
            int myoff = base + incl - mycount;
            __syncthreads();                                  // every count has been read
#pragma unroll
            for (int q = 0; q < kClsPerThread; ++q) {
                hist_s[tid * kClsPerThread + q] = myoff;
                myoff += cnt[q];
            }
            __syncthreads();
            for (int i = tid; i < K; i += kSortThreads) {
                const float4 t = reinterpret_cast<const float4*>(rows + (size_t)i * YV3_CAND_COLS)[1];
                const unsigned long long key = make_key(t.z, t.w, i);
                bkey[hist_s[(int)(key >> 50)] + pos_s[i]] = key;
            }
            __syncthreads();
            unsigned char* keep_b = keep + (size_t)n * capacity;
            // warp w serves the classes w, w + 32, ...; count(c) = offset(c + 1) - offset(c)
            const int cend = maxc_s + 1;                     // (80 of the 1024 class slots at the benchmark shapes)
            for (int c = warp; c < cend; c += kSortThreads / 32) {
                const int o0 = hist_s[c];
                const int o1 = (c + 1 < kBucketClasses) ? hist_s[c + 1] : K;
                const int len = o1 - o0;
                if (len <= 0) continue;
                if (len > kFuseMaxSeg) {
                    // a bucket of 33..256: published as it is (keys of one class, arrival order) and queued with its
                    // length and the "not sorted yet" flag: (image, first position | length << 18 | 1 << 27).  The warp
                    // that takes it off the queue in nms_warp_kernel rank-sorts it in its shared memory -- 20 000
                    // buckets spread over every SM instead of 80 per image serialised on this CTA's 32 warps, where
                    // the rank sort was 79 % of this kernel's instructions at the dense stress workload.
                    if (bkey != gk)
                        for (int p = lane; p < len; p += 32) gk[o0 + p] = bkey[o0 + p];
                    if (lane == 0) wl[atomicAdd(wl_count, 1)] = make_int2(n, o0 | (len << 18) | (1 << 27));
                    continue;
                }
                // rank sort of at most 32 keys, one per lane: keys are unique (the row is part of the key)
                {
                    const unsigned long long mykey = (lane < len) ? bkey[o0 + lane] : kPadKey;
                    int rank = 0;
                    for (int j = 0; j < len; ++j) {
                        const unsigned long long other = __shfl_sync(0xffffffffu, mykey, j);
                        rank += (other < mykey) ? 1 : 0;
                    }
                    __syncwarp();
                    if (lane < len) bkey[o0 + rank] = mykey;
                    __syncwarp();
                }
                const bool in = lane < len;
                const unsigned long long k1 = in ? bkey[o0 + lane] : kPadKey;
                const int row = (int)(k1 & 0x3ffffULL);
                float4 mine_f = make_float4(0.f, 0.f, 0.f, 0.f);
                if (in) {
                    mine_f = reinterpret_cast<const float4*>(rows + (size_t)row * YV3_CAND_COLS)[0];
                }
                unsigned alive = __ballot_sync(0xffffffffu, in);
                const float am = __fmul_rn(__fsub_rn(mine_f.z, mine_f.x), __fsub_rn(mine_f.w, mine_f.y));
                const bool sane = !in || (am >= kWarpNmsAreaLo && am <= 1e29f);
                if (nq.lo >= kWarpNmsThreshLo && __all_sync(0xffffffffu, sane)) {
                    // every pair first, branch-free (the sign tests of warp_nms_tile, candidate i by shuffles): per lane j the
                    // bits i < j that suppress it and the bits the fp32 test leaves open; the latter by the exact expression;
                    // then the survivors by the ballot fixed point x <- in & ~exists(x_i & S_ij) -- 2-3 ballots.  The serial
                    // walk over the kept candidates it replaces spent ~50 instructions per candidate.
                    unsigned col = 0u, und = 0u;
                    for (int i = 0; i < len - 1; ++i) {
                        const float fx = __shfl_sync(0xffffffffu, mine_f.x, i), fy = __shfl_sync(0xffffffffu, mine_f.y, i);
                        const float fz = __shfl_sync(0xffffffffu, mine_f.z, i), fw = __shfl_sync(0xffffffffu, mine_f.w, i);
                        const float ai = __shfl_sync(0xffffffffu, am, i);
                        const float w = fmaxf(__fsub_rn(fminf(fz, mine_f.z), fmaxf(fx, mine_f.x)), 0.f);
                        const float h = __fsub_rn(fminf(fw, mine_f.w), fmaxf(fy, mine_f.y));
                        const float inter = __fmul_rn(w, h);
                        const float uni = __fsub_rn(__fadd_rn(ai, am), inter);
                        const bool yes = __fmaf_rn(nq.hi, uni, -inter) < 0.f;
                        const bool no = __fmaf_rn(-nq.lo, uni, inter) < 0.f;
                        col |= yes ? (1u << i) : 0u;
                        und |= (yes || no) ? 0u : (1u << i);
                    }
                    const unsigned earlier = in ? (1u << lane) - 1u : 0u;               // only EARLIER candidates suppress
                    col &= earlier;
                    und &= earlier;
                    for (unsigned open = __reduce_or_sync(0xffffffffu, und); open; open &= open - 1u) {
                        const int i = __ffs(open) - 1;
                        float4 fi;
                        fi.x = __shfl_sync(0xffffffffu, mine_f.x, i); fi.y = __shfl_sync(0xffffffffu, mine_f.y, i);
                        fi.z = __shfl_sync(0xffffffffu, mine_f.z, i); fi.w = __shfl_sync(0xffffffffu, mine_f.w, i);
                        if ((und >> i) & 1u) {
                            // (an open pair overlaps: a disjoint one has inter <= 0 and falls on the "no" side)
                            if (suppresses_overlapping(fi, box_area_d(fi), mine_f, box_area_d(mine_f), thresh)) col |= 1u << i;
                        }
                    }
                    for (;;) {
                        const unsigned nk = __ballot_sync(0xffffffffu, in && (col & alive) == 0u);
                        if (nk == alive) break;
                        alive = nk;
                    }
                } else {
                    for (int i = 0; i < len - 1; ++i) {
                        if (!((alive >> i) & 1u)) continue;                              // warp-uniform
                        float4 fi;
                        fi.x = __shfl_sync(0xffffffffu, mine_f.x, i); fi.y = __shfl_sync(0xffffffffu, mine_f.y, i);
                        fi.z = __shfl_sync(0xffffffffu, mine_f.z, i); fi.w = __shfl_sync(0xffffffffu, mine_f.w, i);
                        const bool overlap = fminf(fi.z, mine_f.z) > fmaxf(fi.x, mine_f.x) && fminf(fi.w, mine_f.w) > fmaxf(fi.y, mine_f.y);
                        bool drop = lane > i && ((alive >> lane) & 1u);
                        if (drop) drop = overlap ? suppresses_overlapping_fast(fi, mine_f, nq) : (0.0 > thresh);
                        alive &= ~__ballot_sync(0xffffffffu, drop);
                    }
                }
                if (in) keep_b[row] = (unsigned char)((alive >> lane) & 1u);
            }
            return;
        }
        __syncthreads();                                      // the attempt is abandoned: the general path starts over
    }
    for (int i = tid; i < n2; i += kSortThreads) {
        unsigned long long key = kPadKey;
        if (i < K) {
            const float4 t = reinterpret_cast<const float4*>(rows + (size_t)i * YV3_CAND_COLS)[1];
            key = make_key(t.z, t.w, i);
        }
        keys[i] = key;
    }
    __syncthreads();
    bool done = false;
    if (keys == skeys && K > kSortThreads && 2 * K <= n2s)
        done = bin_sort_keys(keys, K, hist_s, reinterpret_cast<unsigned short*>(hist_s + kBucketClasses), tid);
    if (!done) {
        if (keys == skeys && n2 <= kSortThreads) bitonic_sort_regs(keys, n2, tid);
        else bitonic_sort(keys, n2, tid);
    }
    // every class segment goes to a queue: nms_warp_kernel (up to 256 candidates) or the cluster kernel (longer)
    for (int i = tid; i < K; i += kSortThreads) {
        const unsigned long long key = keys[i];
        const int row = (int)(key & 0x3ffffULL);
        if (keys != gk) gk[i] = key;
        sidx[(size_t)n * capacity + i] = row;
        sbox[(size_t)n * capacity + i] = reinterpret_cast<const float4*>(rows + (size_t)row * YV3_CAND_COLS)[0];
        const bool starts = i == 0 || (keys[i - 1] >> 50) != (key >> 50);
        if (starts) {
            const bool large = i + kNmsMedium < K && (keys[i + kNmsMedium] >> 50) == (key >> 50);
            if (large) wl_large[atomicAdd(wl_count + 1, 1)] = make_int2(n, i);
            else wl[atomicAdd(wl_count, 1)] = make_int2(n, i);           // length 0: the warp kernel finds the end itself
        }
    }
}

// ---- greedy NMS per (class, image) segment ---------------------------------------------------------------
constexpr int kKeptSmem = 2048;

__host__ __device__ inline size_t nms_smem_bytes(int T) {
    return (size_t)kKeptSmem * 16 + (size_t)T * 16 + (size_t)T * (T / 64) * 8 + (size_t)T * 2;
}

// exact test on the fp32 coordinates: do the boxes overlap at all (:139-142)?  (min/max/compare are exact)
__device__ __forceinline__ bool overlaps_f32(const float4& a, const float4& b) {
    return fminf(a.z, b.z) > fmaxf(a.x, b.x) && fminf(a.w, b.w) > fmaxf(a.y, b.y);
}

// Greedy NMS of the queued segments, one CTA per segment at a time, T threads = T candidates per chunk: each chunk is
// first tested against the boxes kept so far (all T candidates in parallel), then resolved internally with a T x T
// suppression bit matrix in shared memory and a sweep by one warp.  Pairs that do not overlap are settled in fp32;
// the fp64 IoU runs for the others.
template <int T>
__global__ void __launch_bounds__(T) nms_kernel(const unsigned long long* __restrict__ gkeys,
                                                const int32_t* __restrict__ count, int count_cols,
                                                int capacity, int cap2, double thresh,
                                                const float4* __restrict__ sbox, const int32_t* __restrict__ sidx,
                                                float4* __restrict__ kspill, unsigned char* __restrict__ keep,
                                                const int2* __restrict__ wl, const int32_t* __restrict__ wl_count) {
    constexpr int W = T / 64;                         // 64-bit words per row of the bit matrix
    extern __shared__ __align__(16) unsigned char nsm[];
    float4* kept_s = reinterpret_cast<float4*>(nsm);                          // [kKeptSmem]
    float4* chunk_s = kept_s + kKeptSmem;                                     // [T]
    unsigned long long* mask_s = reinterpret_cast<unsigned long long*>(chunk_s + T);   // [T][W]
    unsigned char* alive_s = reinterpret_cast<unsigned char*>(mask_s + (size_t)T * W); // [T]
    unsigned char* kept_flag_s = alive_s + T;                                 // [T]
    __shared__ int seg_s[2];
    __shared__ int nkept_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwork = *wl_count;
    const bool none_overlap_drops = 0.0 > thresh;      // a negative threshold suppresses even disjoint boxes (:124)
    const NmsThresh nq = make_nms_thresh(thresh);
  for (int wi = blockIdx.x; wi < nwork; wi += gridDim.x) {
    __syncthreads();                                   // shared state of the previous segment is done with
    const int2 job = wl[wi];
    const int n = job.x;
    int K = 0;
    for (int q = 0; q < count_cols; ++q) K += count[n * count_cols + q];
    K = max(0, min(K, capacity));
    const unsigned long long* keys = gkeys + (size_t)n * cap2;
    if (T == kNmsMedium) {
        // a queued medium segment is at most T long: every thread probes one position, the first class change wins
        if (tid == 0) { seg_s[0] = job.y; seg_s[1] = min(K, job.y + T); nkept_s = 0; }
        __syncthreads();
        const unsigned long long c0 = keys[job.y] >> 50;
        const int p = job.y + 1 + tid;
        if (p < K && (keys[p] >> 50) != c0) atomicMin(&seg_s[1], p);
    } else if (tid == 0) {
        // end of the segment: first sorted position whose class field exceeds this segment's
        const unsigned long long want = (keys[job.y] >> 50) + 1ULL;
        int lo = job.y, hi = K;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if ((keys[mid] >> 50) < want) lo = mid + 1; else hi = mid;
        }
        seg_s[0] = job.y;
        seg_s[1] = lo;
        nkept_s = 0;
    }
    __syncthreads();
    const int s0 = seg_s[0], s1 = seg_s[1];
    const float4* boxes = sbox + (size_t)n * capacity;
    float4* spill = kspill + (size_t)n * capacity + s0;
    const int32_t* rowof = sidx + (size_t)n * capacity;
    unsigned char* keep_n = keep + (size_t)n * capacity;

    for (int c0 = s0; c0 < s1; c0 += T) {
        const int cnt = min(T, s1 - c0);
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
            if (alive) {
                const bool drop = overlaps_f32(kb, mine_f) ? suppresses_overlapping_fast(kb, mine_f, nq)
                                                           : none_overlap_drops;
                if (drop) alive = false;
            }
        }
        alive_s[tid] = alive ? 1 : 0;
        __syncthreads();
        // phase 2: bit matrix inside the chunk: row i, bit j (j > i) <=> i suppresses j.  The triangle is cut into tasks
        // of (row, 32 columns) dealt round-robin to the threads, so the thread of row 0 does not test cnt - 1 pairs
        // alone while the last rows idle; a task writes its own 32-bit half word, no atomics.
        {
            unsigned* mask32 = reinterpret_cast<unsigned*>(mask_s);
            const int hw_n = (cnt + 31) >> 5;                       // half words that hold columns < cnt
            for (int i = tid; i < cnt * 2 * W; i += T) {            // rows' unused half words: zero
                const int hw = i % (2 * W);
                if (hw >= hw_n) mask32[i] = 0u;
            }
            for (int task = tid; task < cnt * hw_n; task += T) {
                const int i = task / hw_n, hw = task - i * hw_n;
                unsigned bits = 0u;
                const int jb = max(i + 1, hw << 5), je = min(cnt, (hw + 1) << 5);
                if (jb < je && alive_s[i]) {
                    const float4 oi = chunk_s[i];
                    const double ai = box_area_d(oi);
                    for (int j = jb; j < je; ++j) {
                        if (!alive_s[j]) continue;
                        const float4 oj = chunk_s[j];
                        const bool drop = overlaps_f32(oi, oj) ? suppresses_overlapping_fast(oi, oj, nq)
                                                               : none_overlap_drops;
                        if (drop) bits |= 1u << (j & 31);
                    }
                }
                mask32[(size_t)i * 2 * W + hw] = bits;
            }
        }
        __syncthreads();
        // sweep: warp 0 walks the chunk in order, 64 rows (one word of the removed set) at a time.  Inside a block the
        // decisions depend only on that word, which every lane carries redundantly (rblk) next to its own word
        // (lane w < W: word w); the rows' words are loaded ahead, independent of the decisions, so the serial chain
        // per row is a bit test and two ORs instead of a shuffle and a dependent shared-memory load.
        if (warp == 0) {
            unsigned long long removed = 0;
            for (int i0 = 0; i0 < cnt; i0 += 64) {
                const int blk = i0 >> 6;
                unsigned long long rblk = __shfl_sync(0xffffffffu, removed, blk);
                unsigned long long kb = 0;
#pragma unroll 8
                for (int b = 0; b < 64; ++b) {
                    const int i = i0 + b;                                        // < T: the rows exist, stale ones are masked
                    const bool a = i < cnt && alive_s[i];
                    const unsigned long long diag = mask_s[(size_t)i * W + blk];
                    const unsigned long long own = (lane < W) ? mask_s[(size_t)i * W + lane] : 0ULL;
                    const bool k = a && !((rblk >> b) & 1ULL);
                    if (k) { rblk |= diag; removed |= own; kb |= 1ULL << b; }
                }
                kept_flag_s[i0 + lane] = (unsigned char)((kb >> lane) & 1ULL);
                kept_flag_s[i0 + 32 + lane] = (unsigned char)((kb >> (lane + 32)) & 1ULL);
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
}

// ---- segments of 33..256 candidates: one WARP per segment ---------------------------------------------------------
// The greedy sweep of _apply_nms (:117-128) inside one warp, no block barrier: the segment's keys are rank-sorted in
// the warp's shared-memory slice when they arrive as a bucket of the counting sort, the sorted boxes and their fp32
// areas are staged there, and the sweep runs block by block (32 candidates in the lanes against the earlier blocks'
// survivors broadcast from shared memory) -- see the kernel body.  History at the dense stress workload (cfg4, 20 480
// segments of ~133): nms_kernel<256> with CTA-wide phases 1150 us; candidates in registers, one ballot per (i, slot)
// 950 us; a column bit matrix in shared memory + an 8-lane sweep 690 us; this form (sort included) ~400 us.
constexpr int kWarpNmsThreads = 32;      // host launch: one warp per CTA, 5 KB of shared memory -- an idle launch (nothing
                                         // queued) slips in next to the sweep's CTAs
constexpr int kWarpNmsWarps = kWarpNmsThreads / 32;
// launched from the device (sort_kernel's dispatch) a grid gets the smallest shared-memory carve-out that holds ONE of its
// CTAs, i.e. one CTA per SM (measured with the 14 KB form: 32-thread CTAs 7.9 ms, 128-thread CTAs 2.2 ms at cfg4 against
// 1.1 ms from the host): the device-side launch therefore uses one many-warp CTA per SM
constexpr int kWarpNmsDevThreads = 768;      // 24 warps: 76 registers per thread, 120 KB of shared memory
struct WarpNmsSmem {
    float4 box[kWarpNmsMax];       // the segment's boxes in sorted order (before that: the key buffer of the rank sort)
    float area[kWarpNmsMax];
};
static_assert(sizeof(float4) * kWarpNmsMax >= 2 * sizeof(unsigned long long) * kWarpNmsMax, "the box array doubles as the two key buffers");

// One tile of pair tests: candidate j (this lane: box bj, area aj) against the `iend` candidates i at bi / ai_p.
// Returns in `drop` the bits i whose fp32 test says IoU(i, j) > thr and in `und` the bits the fp32 test does not settle
// (to be resolved by the exact expression); both still unmasked.  The fast loop has no branch and no predicate, 14
// arithmetic instructions per 32 pair tests: with w clamped at 0 a pair that does not overlap has inter <= 0 and falls
// on the "no" side by itself, so per (i, j) two signs are kept, shifted into two words:
//   yes: fma(hi, uni, -inter) < 0  <=>  inter > hi * uni  (the product exact: tighter than the rounded one)
//   no:  fma(-lo, uni, inter) < 0  <=>  inter < lo * uni
__device__ __forceinline__ void warp_nms_tile(const bool fast, const float4* bi, const float* ai_p, const int iend,
                                              const float4 bj, const float aj, const NmsThresh& nq,
                                              const bool none_overlap_drops, unsigned& drop, unsigned& und) {
    drop = 0u;
    und = 0u;
    if (fast) {
        unsigned yrev = 0u, nrev = 0u;
#pragma unroll 4
        for (int l = 0; l < iend; ++l) {
            const float4 fi = bi[l];
            const float ai = ai_p[l];
            const float w = fmaxf(__fsub_rn(fminf(fi.z, bj.z), fmaxf(fi.x, bj.x)), 0.f);
            const float h = __fsub_rn(fminf(fi.w, bj.w), fmaxf(fi.y, bj.y));
            const float inter = __fmul_rn(w, h);
            const float uni = __fsub_rn(__fadd_rn(ai, aj), inter);
            yrev = __funnelshift_l(__float_as_uint(__fmaf_rn(nq.hi, uni, -inter)), yrev, 1);
            nrev = __funnelshift_l(__float_as_uint(__fmaf_rn(-nq.lo, uni, inter)), nrev, 1);
        }
        // iteration l left its sign at bit iend - 1 - l
        const unsigned valid = iend >= 32 ? 0xffffffffu : (1u << iend) - 1u;
        const unsigned yes = __brev(yrev) >> (32 - iend), no = __brev(nrev) >> (32 - iend);
        drop = yes & valid;
        und = ~(yes | no) & valid;
    } else {
        for (int l = 0; l < iend; ++l) {
            const float4 fi = bi[l];
            const bool ov = fminf(fi.z, bj.z) > fmaxf(fi.x, bj.x) && fminf(fi.w, bj.w) > fmaxf(fi.y, bj.y);   // :139-142
            if (ov) und |= 1u << l;
            else if (none_overlap_drops) drop |= 1u << l;               // a negative threshold: disjoint boxes suppress too (:124)
        }
    }
}

// Sort of a bucket of `len` <= 256 keys (arrival order -> ascending key = score descending, row ascending) by one warp:
// every slot of 32 keys is sorted in the lanes by a bitonic network of shuffles (15 compare-exchange rounds) and put down
// in `skey`; a key's final rank is then its place in its own slot plus, for every other slot, the number of keys there
// below it -- a binary search over that sorted slot (keys are unique: the row is part of the key; the padding of the last
// slot is the largest key and sorts to its end).  The sorted bucket is written to `sorted` (another buffer: the slots are
// still being searched).  ~1.3 K instructions for 133 keys; counting every key against every other one took 2.3 K.
__device__ __forceinline__ void warp_sort_bucket(unsigned long long* skey, unsigned long long* sorted, const int lane,
                                                 const int len, const int nslot, const unsigned long long* keys) {
    for (int u = 0; u < nslot; ++u) {
        unsigned long long k = (lane + 32 * u < len) ? keys[lane + 32 * u] : kPadKey;
#pragma unroll
        for (int kk = 2; kk <= 32; kk <<= 1) {
#pragma unroll
            for (int j = kk >> 1; j > 0; j >>= 1) {
                const unsigned long long o = __shfl_xor_sync(0xffffffffu, k, j);
                const bool take_min = ((lane & j) == 0) == ((lane & kk) == 0);       // (lane & 32) == 0: the last merge ascends
                const bool less = k < o;
                k = (take_min == less) ? k : o;
            }
        }
        skey[32 * u + lane] = k;
    }
    __syncwarp();
    for (int u = 0; u < nslot; ++u) {
        const unsigned long long k = skey[32 * u + lane];
        int rank = lane;
        for (int w = 0; w < nslot; ++w) {
            if (w == u) continue;
            const unsigned long long* slot = skey + 32 * w;
            int lo = 0;
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1)
                if (slot[lo + step - 1] < k) lo += step;
            lo += (slot[lo] < k) ? 1 : 0;                           // lo <= 31 here; elements [0, lo) are below k
            rank += lo;
        }
        if (k != kPadKey) sorted[rank] = k;
    }
    __syncwarp();
}

// One queued segment, by one warp (`sm`: the warp's slice of shared memory).  Called by nms_warp_kernel and, for a short
// queue, by the last CTA of the sort kernel itself.
__device__ __forceinline__ void warp_nms_job(WarpNmsSmem& sm, const int lane, const int2 job,
                                             const unsigned long long* __restrict__ gkeys, const int32_t* __restrict__ count,
                                             const int count_cols, const int capacity, const int cap2, const double thresh,
                                             const NmsThresh nq, const bool fast_ok, const bool none_overlap_drops,
                                             const float* __restrict__ cand, unsigned char* __restrict__ keep) {
    {
        const int n = job.x, s0 = job.y & 0x3ffff;
        int len = (job.y >> 18) & 0x1ff;
        const bool needs_sort = ((job.y >> 27) & 1) != 0;
        const unsigned long long* keys = gkeys + (size_t)n * cap2 + s0;
        if (len == 0) {
            // queued by the general sort path without a length: the run of this class from s0 on (at most 256 long)
            int K = 0;
            for (int q = 0; q < count_cols; ++q) K += count[n * count_cols + q];
            K = max(0, min(K, capacity));
            const unsigned long long c0 = keys[0] >> 50;
#pragma unroll
            for (int u = 0; u < kWarpNmsSlots; ++u) {
                const int p = s0 + lane + 32 * u;
                len += __popc(__ballot_sync(0xffffffffu, p < K && (keys[lane + 32 * u] >> 50) == c0));
            }
        }
        const float* rows = cand + (size_t)n * capacity * YV3_CAND_COLS;
        const int nslot = (len + 31) >> 5;
        __syncwarp();                                                    // the previous segment's sweep is done with sm
        if (needs_sort) {
            // the box array is free until the boxes are staged: its halves are the slot buffer and the sorted bucket
            unsigned long long* skey = reinterpret_cast<unsigned long long*>(&sm.box[0]);
            warp_sort_bucket(skey, skey + kWarpNmsMax, lane, len, nslot, keys);
            skey += kWarpNmsMax;
            keys = skey;
        }
        int row[kWarpNmsSlots];
#pragma unroll
        for (int u = 0; u < kWarpNmsSlots; ++u) row[u] = (lane + 32 * u < len) ? (int)(keys[lane + 32 * u] & 0x3ffffULL) : -1;
        __syncwarp();                                                    // the keys are read: their buffer becomes the box array
        bool sane = true;
#pragma unroll
        for (int u = 0; u < kWarpNmsSlots; ++u) {
            if (row[u] >= 0) {
                const float4 b = __ldg(reinterpret_cast<const float4*>(rows + (size_t)row[u] * YV3_CAND_COLS));
                const float a = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
                sm.box[lane + 32 * u] = b;
                sm.area[lane + 32 * u] = a;
                sane = sane && a >= kWarpNmsAreaLo && a <= 1e29f;
            }
        }
        // one candidate outside that range (or a threshold below 1e-6) sends the segment to the exact expression pair by
        // pair: correct, slow, and not met in practice
        const bool fast = fast_ok && __all_sync(0xffffffffu, sane);
        __syncwarp();
        // ---- the greedy sweep of _apply_nms (:117-128), block by block.  Blocks are cut so that the RAGGED one comes
        // first: block 0 = the first r = len - 32 (nslot - 1) candidates, blocks 1.. = 32 each (the ragged block is then
        // the one every later block meets, r iterations each, instead of the one that meets every block).  Per block b:
        //   1. diagonal tile: lane j tests its candidate against the block's earlier ones (bits below its own);
        //   2. the block's survivors: x_j = alive_j and no kept i < j suppresses j.  Iterating
        //      x <- alive & ~exists(x_i & S_ij) from x = alive fixes lane t after t rounds (induction on j) and stops at
        //      the unique solution, usually in 2-3 ballots instead of a 32-step chain;
        //   3. every later block tests its candidates against block b's SURVIVORS only (all of its members when most
        //      survived, the survivors one by one when few did) and marks the suppressed ones dead.
        // No suppression matrix is kept: 5 KB of shared memory per warp instead of 14, so the warps of this kernel and
        // the CTAs of the sweep that runs next to it both fit on an SM.
        const int r = len - 32 * (nslot - 1);                            // 1..32
        unsigned deadmask = 0u;                                          // bit v: this lane's candidate of block v is suppressed
        unsigned keptw = 0u;                                             // lane b: the survivors of block b
        for (int ib = 0; ib < nslot; ++ib) {
            const int ibase = ib == 0 ? 0 : r + 32 * (ib - 1);
            const int iend = ib == 0 ? r : 32;
            const float4* bi = sm.box + ibase;
            const float* ai_p = sm.area + ibase;
            unsigned km;
            {
                const bool inj = lane < iend;
                const float4 bj = inj ? bi[lane] : make_float4(0.f, 0.f, 0.f, 0.f);
                const float aj = inj ? ai_p[lane] : 1.f;
                const bool alive = inj && !((deadmask >> ib) & 1u);
                unsigned drop = 0u, und = 0u;
                if (__any_sync(0xffffffffu, alive)) {
                    warp_nms_tile(fast, bi, ai_p, iend, bj, aj, nq, none_overlap_drops, drop, und);
                    const unsigned allow = alive ? (1u << lane) - 1u : 0u;          // only EARLIER candidates suppress
                    drop &= allow;
                    und &= allow;
                    while (und) {
                        const int l = __ffs(und) - 1;
                        und &= und - 1u;
                        const float4 fi = bi[l];
                        if (suppresses_overlapping(fi, box_area_d(fi), bj, box_area_d(bj), thresh)) drop |= 1u << l;
                    }
                }
                km = __ballot_sync(0xffffffffu, alive);
                for (;;) {
                    const unsigned nk = __ballot_sync(0xffffffffu, alive && (drop & km) == 0u);
                    if (nk == km) break;
                    km = nk;
                }
                if (lane == ib) keptw = km;
            }
            if (km == 0u) continue;
            const bool sparse = __popc(km) * 4 < iend * 3;               // fewer than 3/4 survived: visit the survivors only
            for (int v = ib + 1; v < nslot; ++v) {
                const int jbase = r + 32 * (v - 1);                      // blocks 1.. are full: every lane holds a candidate
                const bool alive = !((deadmask >> v) & 1u);
                if (!__any_sync(0xffffffffu, alive)) continue;
                const float4 bj = sm.box[jbase + lane];
                const float aj = sm.area[jbase + lane];
                unsigned drop = 0u, und = 0u;
                if (sparse && fast) {
                    unsigned ysign = 0u, decided = 0xffffffffu;
                    for (unsigned m = km; m; m &= m - 1u) {
                        const int l = __ffs(m) - 1;
                        const float4 fi = bi[l];
                        const float ai = ai_p[l];
                        const float w = fmaxf(__fsub_rn(fminf(fi.z, bj.z), fmaxf(fi.x, bj.x)), 0.f);
                        const float h = __fsub_rn(fminf(fi.w, bj.w), fmaxf(fi.y, bj.y));
                        const float inter = __fmul_rn(w, h);
                        const float uni = __fsub_rn(__fadd_rn(ai, aj), inter);
                        const unsigned y = __float_as_uint(__fmaf_rn(nq.hi, uni, -inter));
                        const unsigned o = __float_as_uint(__fmaf_rn(-nq.lo, uni, inter));
                        ysign |= y;
                        decided &= y | o;
                    }
                    if (ysign >> 31) drop = km;                          // some survivor suppresses it: which one is irrelevant
                    else if (!(decided >> 31)) und = km;                 // a pair inside the band: the exact walk below (rare)
                    if (und) {                                           // not every bit is undecided: only overlapping pairs go on
                        unsigned uo = 0u;
                        for (unsigned m = km; m; m &= m - 1u) {
                            const int l = __ffs(m) - 1;
                            const float4 fi = bi[l];
                            if (fminf(fi.z, bj.z) > fmaxf(fi.x, bj.x) && fminf(fi.w, bj.w) > fmaxf(fi.y, bj.y)) uo |= 1u << l;
                        }
                        und = uo;
                    }
                } else {
                    warp_nms_tile(fast, bi, ai_p, iend, bj, aj, nq, none_overlap_drops, drop, und);
                    drop &= km;
                    und &= km;
                }
                bool hit = drop != 0u;
                if (alive && !hit) {
                    while (und) {
                        const int l = __ffs(und) - 1;
                        und &= und - 1u;
                        const float4 fi = bi[l];
                        if (suppresses_overlapping(fi, box_area_d(fi), bj, box_area_d(bj), thresh)) { hit = true; break; }
                    }
                }
                if (hit) deadmask |= 1u << v;
            }
        }
        unsigned char* keep_n = keep + (size_t)n * capacity;
#pragma unroll
        for (int u = 0; u < kWarpNmsSlots; ++u) {
            const int idx = lane + 32 * u;
            const int b = idx < r ? 0 : 1 + ((idx - r) >> 5);
            const int pos = idx < r ? idx : (idx - r) & 31;
            const unsigned k = __shfl_sync(0xffffffffu, keptw, b & 31);
            if (row[u] >= 0) keep_n[row[u]] = (unsigned char)((k >> pos) & 1u);
        }
    }
}


__global__ void __launch_bounds__(kWarpNmsDevThreads) nms_warp_kernel(const unsigned long long* __restrict__ gkeys,
                                                                   const int32_t* __restrict__ count, int count_cols,
                                                                   int capacity, int cap2, double thresh,
                                                                   const float* __restrict__ cand,
                                                                   unsigned char* __restrict__ keep,
                                                                   const int2* __restrict__ wl, const int32_t* __restrict__ wl_count) {
    extern __shared__ __align__(16) unsigned char wsm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpNmsSmem& sm = reinterpret_cast<WarpNmsSmem*>(wsm)[warp];
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nw = (gridDim.x * blockDim.x) >> 5;
    const int nwork = *wl_count;
    const bool none_overlap_drops = 0.0 > thresh;      // a negative threshold suppresses even disjoint boxes (:124)
    const NmsThresh nq = make_nms_thresh(thresh);
    const bool fast_ok = nq.lo >= kWarpNmsThreshLo;
    for (int wi = wid; wi < nwork; wi += nw)
        warp_nms_job(sm, lane, wl[wi], gkeys, count, count_cols, capacity, cap2, thresh, nq, fast_ok, none_overlap_drops, cand, keep);
}

// ---- long segments: one thread-block CLUSTER per segment ------------------------------------------------------
// A segment of thousands of candidates (a single-class head with a low threshold) is a serial chain of 1024-candidate
// chunks, and inside a chunk the work is (candidates x boxes kept so far) + (chunk x chunk) pair tests: one CTA took
// milliseconds per image.  Here the CTAs of a cluster share it.  Per chunk:
//   1. every CTA holds the chunk (one candidate per thread) and tests it against ITS share of the kept list (kept box q
//      belongs to CTA q mod kClu), staged through shared memory with the fp32 areas, with the branch-free fused sign
//      test; the verdicts of the CTAs are AND-ed through distributed shared memory;
//   2. the candidates still alive (typically a third of the chunk) are compacted, in order, by every CTA alike; the
//      chunk's own suppression matrix is computed on the compact list as 32 x 32 warp tiles dealt over all warps of the
//      cluster and written, row form, straight into the leader's shared memory (diagonal tiles also in column form);
//   3. one warp of the leader sweeps the compact list block by block -- a block's survivors by the ballot fixed point,
//      their rows OR-ed into the later blocks by one warp reduction per block -- appends the survivors to the kept list
//      (global memory, L2-resident) and publishes the count.
// Three cluster barriers per chunk; the greedy order -- hence the result -- is unchanged.  What is left is phase 1
// (issue-bound: 1024 x kept / kClu pair tests per CTA and chunk, ~60 % of the kernel at the single-class stress shape).
// CTAs per cluster.  Measured on B200 at cfg5 (16 single-class segments of 8 053 candidates) with the first version of
// this kernel: 8 -> 2.17 ms (15 clusters fit: two rounds), 7 -> 2.56, 6 -> 1.44 (one round), 4 -> 1.86 ms per step; 8 is
// the portable maximum.  This version: 0.35 ms for the kernel, 0.48 ms per step.
#ifndef YV3_NMS_CLUSTER
#define YV3_NMS_CLUSTER 6
#endif
constexpr int kClu = YV3_NMS_CLUSTER;
constexpr int kCluT = 1024;             // threads per CTA = candidates per chunk
constexpr int kCluRow = kCluT / 32 + 1;  // words per row of the chunk's suppression matrix: 32 + 1, so that a column
                                        // (one word of 32 consecutive rows) is read without a bank conflict
__host__ __device__ inline size_t nms_cluster_smem_bytes() {
    return (size_t)kCluT * 16 + (size_t)kCluT * kCluRow * 4 + (size_t)kCluT * 3 + (size_t)kCluT * 20 + (size_t)kCluT * 6;
}

__global__ void __launch_bounds__(kCluT, 1) nms_cluster_kernel(const unsigned long long* __restrict__ gkeys,
                                                               const int32_t* __restrict__ count, int count_cols,
                                                               int capacity, int cap2, double thresh,
                                                               const float4* __restrict__ sbox, const int32_t* __restrict__ sidx,
                                                               float4* __restrict__ kspill, unsigned char* __restrict__ keep,
                                                               const int2* __restrict__ wl, const int32_t* __restrict__ wl_count) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    constexpr int T = kCluT;
    extern __shared__ __align__(16) unsigned char csm[];
    float4* chunk_s = reinterpret_cast<float4*>(csm);                                         // [T]
    unsigned* mask_s = reinterpret_cast<unsigned*>(chunk_s + T);                             // [T][kCluRow] row form (used in the leader)
    unsigned char* verdict_s = reinterpret_cast<unsigned char*>(mask_s + (size_t)T * kCluRow);   // [T] this CTA's share
    unsigned char* spare_s = verdict_s + T;                                                  // [T] (free)
    unsigned char* kept_flag_s = spare_s + T;                                                // [T]
    float4* kbox_s = reinterpret_cast<float4*>(kept_flag_s + T);                            // [T] a tile of this CTA's share of the kept list
    float* karea_s = reinterpret_cast<float*>(kbox_s + T);                                  // [T]
    unsigned* dcol_s = reinterpret_cast<unsigned*>(karea_s + T);                            // [T] diagonal tiles, column form (leader)
    unsigned short* cidx_s = reinterpret_cast<unsigned short*>(dcol_s + T);                 // [T] chunk position of compact candidate k
    __shared__ int seg_s[2];
    __shared__ int sane_s;
    __shared__ int wcount_s[T / 32];
    __shared__ int na_s;
    __shared__ int nkept_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int r = (int)cluster.block_rank();
    const int cid = blockIdx.x / kClu, ncl = gridDim.x / kClu;
    unsigned* leader_mask32 = cluster.map_shared_rank(mask_s, 0);
    const int* leader_nkept = cluster.map_shared_rank(&nkept_s, 0);
    const int nwork = *wl_count;
    const bool none_overlap_drops = 0.0 > thresh;
    const NmsThresh nq = make_nms_thresh(thresh);
    for (int wi = cid; wi < nwork; wi += ncl) {
        const int2 job = wl[wi];
        const int n = job.x;
        int K = 0;
        for (int q = 0; q < count_cols; ++q) K += count[n * count_cols + q];
        K = max(0, min(K, capacity));
        const unsigned long long* keys = gkeys + (size_t)n * cap2;
        if (tid == 0) {
            // end of the segment: first sorted position whose class field exceeds this segment's (every CTA finds it itself)
            const unsigned long long want = (keys[job.y] >> 50) + 1ULL;
            int lo = job.y, hi = K;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if ((keys[mid] >> 50) < want) lo = mid + 1; else hi = mid;
            }
            seg_s[0] = job.y;
            seg_s[1] = lo;
            nkept_s = 0;
        }
        __syncthreads();
        const int s0 = seg_s[0], s1 = seg_s[1];
        const float4* boxes = sbox + (size_t)n * capacity;
        float4* spill = kspill + (size_t)n * capacity + s0;          // the kept list of this segment, in order
        const int32_t* rowof = sidx + (size_t)n * capacity;
        unsigned char* keep_n = keep + (size_t)n * capacity;
        int nk = 0;
        for (int c0 = s0; c0 < s1; c0 += T) {
            const int cnt = min(T, s1 - c0);
#ifdef YV3_CLUSTER_TIMING
            long long tq[8];
            const bool tq_on = blockIdx.x == 0 && tid == 0;
#define YV3_TQ(k) do { if (tq_on) tq[k] = clock64(); } while (0)
            YV3_TQ(0);
#else
#define YV3_TQ(k) do { } while (0)
#endif
            // phase 1: this thread's candidate against this CTA's share of the boxes kept so far (kept box q belongs to CTA
            // q mod kClu).  The share is staged through shared memory a tile of T boxes at a time, with the fp32 areas, and
            // every warp walks the tile with the branch-free test of warp_nms_tile: two fused sign tests per pair, OR-ed
            // (which kept box suppresses the candidate is irrelevant).  Only a candidate with a pair inside the +-2^-18
            // band -- or a list with an area outside the trusted range -- walks its share again with the exact test.
            bool alive = tid < cnt;
            float4 mine_f = make_float4(0.f, 0.f, 0.f, 0.f);
            float aj = 1.f;
            if (alive) {
                mine_f = boxes[c0 + tid];
                chunk_s[tid] = mine_f;
                aj = __fmul_rn(__fsub_rn(mine_f.z, mine_f.x), __fsub_rn(mine_f.w, mine_f.y));
            }
            const bool sane_mine = aj >= kWarpNmsAreaLo && aj <= 1e29f;
            unsigned ysign = 0u, decided = 0xffffffffu;
            if (tid == 0) sane_s = 1;
            for (int base = 0; base < nk; base += kClu * T) {
                __syncthreads();                                     // the previous tile is consumed (and sane_s is set)
                const int q = base + r + kClu * tid;
                if (q < nk) {
                    const float4 b = __ldcg(spill + q);
                    const float a = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
                    kbox_s[tid] = b;
                    karea_s[tid] = a;
                    if (!(a >= kWarpNmsAreaLo && a <= 1e29f)) sane_s = 0;
                }
                const int rem = nk - base - r;
                const int cntk = rem <= 0 ? 0 : min(T, (rem + kClu - 1) / kClu);
                __syncthreads();
#pragma unroll 4
                for (int l = 0; l < cntk; ++l) {
                    const float4 fi = kbox_s[l];
                    const float ai = karea_s[l];
                    const float w = fmaxf(__fsub_rn(fminf(fi.z, mine_f.z), fmaxf(fi.x, mine_f.x)), 0.f);
                    const float h = __fsub_rn(fminf(fi.w, mine_f.w), fmaxf(fi.y, mine_f.y));
                    const float inter = __fmul_rn(w, h);
                    const float uni = __fsub_rn(__fadd_rn(ai, aj), inter);
                    const unsigned y = __float_as_uint(__fmaf_rn(nq.hi, uni, -inter));
                    const unsigned o = __float_as_uint(__fmaf_rn(-nq.lo, uni, inter));
                    ysign |= y;
                    decided &= y | o;
                }
            }
            __syncthreads();
            {
                const bool fast = nq.lo >= kWarpNmsThreshLo && sane_s != 0 && sane_mine;
                bool hit = fast && (ysign >> 31) != 0u;
                if (alive && !hit && !(fast && (decided >> 31) != 0u)) {
                    for (int q = r; q < nk; q += kClu) {
                        const float4 kb = __ldcg(spill + q);
                        const bool drop = overlaps_f32(kb, mine_f) ? suppresses_overlapping_fast(kb, mine_f, nq) : none_overlap_drops;
                        if (drop) { hit = true; break; }
                    }
                }
                if (hit) alive = false;
            }
            verdict_s[tid] = alive ? 1 : 0;
            YV3_TQ(1);
            cluster.sync();                                          // (A) every CTA's verdicts are in place
            YV3_TQ(2);
            bool all;
            {
                all = tid < cnt;
#pragma unroll
                for (int k = 0; k < kClu; ++k) {
                    const unsigned char* rem = cluster.map_shared_rank(verdict_s, k);
                    all = all && (rem[tid] != 0);
                }
            }
            // ---- the candidates the kept list left alive, compacted in order (every CTA builds the same list): typically a
            // third of the chunk, so the chunk's own pair tests and the sweep shrink ninefold.  cbox / carea reuse the
            // kept-list tile of phase 1.
            float4* cbox_s = kbox_s;
            float* carea_s = karea_s;
            {
                const unsigned bal = __ballot_sync(0xffffffffu, all);
                if (lane == 0) wcount_s[warp] = __popc(bal);
                __syncthreads();                                     // (also: phase 1's last tile is consumed everywhere)
                int base = 0;
                for (int w2 = 0; w2 < warp; ++w2) base += wcount_s[w2];
                if (tid == T - 1) na_s = base + __popc(bal);
                if (all) {
                    const int pos = base + __popc(bal & ((1u << lane) - 1u));
                    cbox_s[pos] = mine_f;
                    carea_s[pos] = aj;
                    cidx_s[pos] = (unsigned short)tid;
                }
                kept_flag_s[tid] = 0;
                __syncthreads();
            }
            const int na = na_s;
            const int nb = (na + 31) >> 5;
            YV3_TQ(3);
            const bool fast2 = nq.lo >= kWarpNmsThreshLo && sane_s != 0;       // (a candidate with an area out of range died... or not:
                                                                                //  checked per tile below)
            // phase 2: the chunk's own suppression matrix over the compact list, one 32 x 32 tile (block ib of earlier
            // candidates i, block v >= ib of later ones j) per warp and turn, dealt over the cluster's warps.  Lanes hold
            // the j's; candidate i is a broadcast read; "i suppresses j" is one fused sign test per pair and a ballot turns
            // the 32 verdicts into row i's word for block v, written straight into the leader's matrix (row form: the
            // sweep ORs the rows of kept candidates).  The diagonal tiles also leave their column form (per j: the earlier
            // i of its own block that suppress it) for the leader's in-block fixed point.
            unsigned* leader_dcol = cluster.map_shared_rank(dcol_s, 0);
            {
                const int ntile = nb * (nb + 1) / 2;
                for (int t = warp * kClu + r; t < ntile; t += kClu * (T / 32)) {       // tile t -> CTA t mod kClu: all CTAs busy
                    int v = (int)((sqrtf(8.f * (float)t + 1.f) - 1.f) * 0.5f);
                    while (v * (v + 1) / 2 > t) --v;
                    while ((v + 1) * (v + 2) / 2 <= t) ++v;
                    const int ib = t - v * (v + 1) / 2;
                    const int j = 32 * v + lane;
                    const bool inj = j < na;
                    const float4 bj = inj ? cbox_s[j] : make_float4(0.f, 0.f, 0.f, 0.f);
                    const float ajj = inj ? carea_s[j] : 1.f;
                    const int iend = min(32, na - 32 * ib);
                    // column form first (lane j: the bits i of block ib that suppress j), with the pipelined sign-word loop
                    // of warp_nms_tile and the exact expression for what it leaves open; a block with an area outside the
                    // trusted range goes pair by pair
                    bool sane = !inj || (ajj >= kWarpNmsAreaLo && ajj <= 1e29f);
                    if (lane < iend) {
                        const float ai = carea_s[32 * ib + lane];
                        sane = sane && ai >= kWarpNmsAreaLo && ai <= 1e29f;
                    }
                    const bool fast = fast2 && __all_sync(0xffffffffu, sane);
                    unsigned col, und;
                    warp_nms_tile(fast, cbox_s + 32 * ib, carea_s + 32 * ib, iend, bj, ajj, nq, none_overlap_drops, col, und);
                    unsigned allow = 0u;                                                   // only EARLIER candidates suppress
                    if (inj) allow = (ib == v) ? (1u << lane) - 1u : 0xffffffffu;
                    col &= allow;
                    und &= allow;
                    while (und) {
                        const int l = __ffs(und) - 1;
                        und &= und - 1u;
                        const float4 fi = cbox_s[32 * ib + l];
                        if (suppresses_overlapping(fi, box_area_d(fi), bj, box_area_d(bj), thresh)) col |= 1u << l;
                    }
                    // ... then row form by ballots: row i's word for block v has bit j set iff i suppresses j
                    unsigned myrow = 0u;
#pragma unroll 8
                    for (int l = 0; l < 32; ++l) {
                        const unsigned bits = __ballot_sync(0xffffffffu, (col >> l) & 1u);
                        if (lane == l) myrow = bits;
                    }
                    if (lane < iend) leader_mask32[(size_t)(32 * ib + lane) * kCluRow + v] = myrow;
                    if (ib == v && inj) leader_dcol[j] = col;
                }
            }
            YV3_TQ(4);
            cluster.sync();                                          // (B) the leader's matrix is complete
            YV3_TQ(5);
            if (r == 0) {
                if (warp == 0) {
                    // phase 3, the greedy sweep over the compact list, block by block: lane w holds the removed bits of
                    // block w; a block's survivors x_j = free_j and no kept i < j of the block suppresses j come from
                    // iterating x <- free & ~exists(x_i & S_ij) (fixed after at most 32 rounds, usually 2-3 ballots); the
                    // rows of the survivors are then OR-ed into the removed bits of the later blocks.
                    const unsigned* m32 = mask_s;
                    unsigned removed = 0u, keptw = 0u;
                    for (int b2 = 0; b2 < nb; ++b2) {
                        const unsigned rem_b = __shfl_sync(0xffffffffu, removed, b2);
                        const int j = 32 * b2 + lane;
                        const bool fr = j < na && !((rem_b >> lane) & 1u);
                        const unsigned dword = (j < na) ? dcol_s[j] : 0u;
                        unsigned km = __ballot_sync(0xffffffffu, fr);
                        for (;;) {
                            const unsigned nkm = __ballot_sync(0xffffffffu, fr && (dword & km) == 0u);
                            if (nkm == km) break;
                            km = nkm;
                        }
                        if (lane == b2) keptw = km;
                        // lane i of the block, if it survived, holds row i; word w of the rows, OR-ed over the lanes, is
                        // what block b2's survivors remove from block w
                        const unsigned* myrowp = m32 + (size_t)(32 * b2 + lane) * kCluRow;
                        const bool mine_kept = (km >> lane) & 1u;
                        for (int w = b2 + 1; w < nb; ++w) {
                            const unsigned o = __reduce_or_sync(0xffffffffu, mine_kept ? myrowp[w] : 0u);
                            if (lane == w) removed |= o;
                        }
                    }
                    // survivors -> flags at their chunk positions, and appended to the kept list in order
                    int total = nk;
                    for (int b2 = 0; b2 < nb; ++b2) {
                        const unsigned km = __shfl_sync(0xffffffffu, keptw, b2);
                        const int k2 = 32 * b2 + lane;
                        const bool kp = (km >> lane) & 1u;
                        if (kp) {
                            kept_flag_s[cidx_s[k2]] = 1;
                            spill[total + __popc(km & ((1u << lane) - 1u))] = cbox_s[k2];
                        }
                        total += __popc(km);
                    }
                    if (lane == 0) nkept_s = total;
                }
                __syncthreads();
                if (tid < cnt) keep_n[rowof[c0 + tid]] = kept_flag_s[tid];
                __threadfence();                                     // the kept list is visible before the barrier below
            }
            YV3_TQ(6);
            cluster.sync();                                          // (C) kept list and count published
            YV3_TQ(7);
#ifdef YV3_CLUSTER_TIMING
            if (tq_on) printf("chunk %d nk %d na %d: ph1 %lld syncA %lld and+compact %lld ph2 %lld syncB %lld ph3 %lld syncC %lld (cycles)\n",
                              (c0 - s0) / T, nk, na, tq[1] - tq[0], tq[2] - tq[1], tq[3] - tq[2], tq[4] - tq[3], tq[5] - tq[4], tq[6] - tq[5], tq[7] - tq[6]);
#endif
            nk = *leader_nkept;
        }
        __syncthreads();
    }
    cluster.sync();                                                  // nobody leaves while its shared memory may be read
}

// The short-queue service of the sort kernel's last CTA (see sort_kernel).  Not inlined: its registers (the warp NMS needs
// 74) stay out of the sort kernel's own allocation.
__device__ __noinline__ void sort_inline_jobs(WarpNmsSmem* sm_base, const int inline_warps, const int nmed,
                                              const unsigned long long* __restrict__ gkeys, const int32_t* __restrict__ count,
                                              const int count_cols, const int capacity, const int cap2, const double thresh,
                                              const float* __restrict__ cand, unsigned char* __restrict__ keep,
                                              const int2* __restrict__ wl) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp >= inline_warps) return;
    const bool none_overlap_drops = 0.0 > thresh;
    const NmsThresh nq = make_nms_thresh(thresh);
    const bool fast_ok = nq.lo >= kWarpNmsThreshLo;
    for (int wi = warp; wi < nmed; wi += inline_warps)
        warp_nms_job(sm_base[warp], lane, wl[wi], gkeys, count, count_cols, capacity, cap2, thresh, nq, fast_ok, none_overlap_drops,
                     cand, keep);
}

// ---- the sort kernel and its device-side dispatch ----------------------------------------------------------------
// The queue-driven NMS kernels above have nothing to do on most steps of a many-class head (no class of an image holds
// more than 32 candidates), yet launched from the host they still had to be placed: thousands of CTAs on the
// high-priority stream, each holding back the sweep's CTAs while the block scheduler found room -- ~6 us of a 103 us
// step at cfg2 (tools/step_breakdown.py).  With `dispatch` set the LAST sort CTA to finish looks at the two queue
// counters and tail-launches (CUDA dynamic parallelism, cudaStreamTailLaunch: after this grid, before the stream's
// next kernel) only the kernels that have work, with grids sized by the queues.
// MINB = 2: the register cap that lets two sort CTAs share an SM (32 registers per thread; the rarely run parts spill) --
// used when there are more images than SMs, so that the grid is one wave instead of two (cfg4: 256 images).
template <int MINB>
__global__ void __launch_bounds__(kSortThreads, MINB) sort_kernel(const float* __restrict__ cand, const int32_t* __restrict__ count,
                                                            int count_cols, int capacity, int cap2, int n2s, double thresh,
                                                            unsigned long long* __restrict__ gkeys, float4* __restrict__ sbox,
                                                            int32_t* __restrict__ sidx, unsigned char* __restrict__ keep,
                                                            int2* __restrict__ wl, int2* __restrict__ wl_large,
                                                            int32_t* __restrict__ wl_count, float4* __restrict__ kspill,
                                                            int dispatch, int many_classes) {
    sort_body(cand, count, count_cols, capacity, cap2, n2s, thresh, gkeys, sbox, sidx, keep, wl, wl_large, wl_count, many_classes != 0);
    if (!dispatch) return;
    __shared__ int last_s;
    __syncthreads();                                            // (every thread of the CTA reaches this point)
    if (threadIdx.x == 0) {
        __threadfence();
        last_s = (atomicAdd(wl_count + 2, 1) == (int)gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!last_s) return;
    __threadfence();
    const int nmed = *(volatile int32_t*)wl_count, nlarge = *(volatile int32_t*)(wl_count + 1);
    // A SHORT queue of medium segments (a class of one image that happens to exceed 32 candidates: a few per step at 608^2)
    // is served right here by the warps of this last CTA, whose shared memory is free now: a device-side launch costs
    // ~50 us of latency on the post-process chain, more than the work (measured: 178 -> 235 us per step at cfg3 for the
    // ranks whose data had one such class).
    const int inline_warps = min(kSortInlineWarps, (int)(((size_t)n2s * 10 + (size_t)kBucketClasses * 4) / sizeof(WarpNmsSmem)));
    bool launch_med = nmed > 0;
    if (nmed > 0 && inline_warps > 0 && nmed <= kSortInlineJobs) {
        extern __shared__ __align__(16) unsigned long long skeys_again[];
        sort_inline_jobs(reinterpret_cast<WarpNmsSmem*>(skeys_again), inline_warps, nmed, gkeys, count, count_cols, capacity, cap2,
                         thresh, cand, keep, wl);
        launch_med = false;
    }
    if (threadIdx.x != 0) return;
    wl_count[2] = 0;
    if (launch_med) {
        const int wpc = dispatch;                                // warps per CTA of the device-launched grid (one CTA per SM)
        nms_warp_kernel<<<min((nmed + wpc - 1) / wpc, 148), 32 * wpc, sizeof(WarpNmsSmem) * wpc, cudaStreamTailLaunch>>>(
            gkeys, count, count_cols, capacity, cap2, thresh, cand, keep, wl, wl_count);
    }
    if (nlarge > 0)
        nms_kernel<kNmsLarge><<<min(nlarge, 148), kNmsLarge, nms_smem_bytes(kNmsLarge), cudaStreamTailLaunch>>>(
            gkeys, count, count_cols, capacity, cap2, thresh, sbox, sidx, kspill, keep, wl_large, wl_count + 1);
}

// ---- kept rows, resized (yolov3_post_process.py:94-104, :162-172) -------------------------------------------
// One CTA per image: the rows apply_nms returns, per head in local-id order, with box[:4] * target_size and the id
// column, as float64 (the reference's rows are float64; an fp32 coordinate times an integer size is exact there).
// bug_compat: a row of head h is returned iff its LOCAL id is the local id of a greedy survivor of ANY head (:81-104,
// start_index never advances); otherwise exactly the survivors.
constexpr int kCollectThreads = 256;
__global__ void __launch_bounds__(kCollectThreads) collect_kernel(const float* __restrict__ cand, const int32_t* __restrict__ counts,
                                                                  const unsigned char* __restrict__ keep, int capacity,
                                                                  double tw, double th, int bug_compat,
                                                                  double* __restrict__ rows_out, int32_t* __restrict__ counts_out) {
    __shared__ int wsum_s[kCollectThreads / 32];
    __shared__ int carry_s;
    const int n = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int c[3], off[4];
    off[0] = 0;
    for (int h = 0; h < 3; ++h) {
        c[h] = max(0, min(counts[n * 3 + h], capacity - off[h]));
        off[h + 1] = off[h] + c[h];
    }
    const float* rows = cand + (size_t)n * capacity * YV3_CAND_COLS;
    const unsigned char* kp = keep + (size_t)n * capacity;
    double* out = rows_out + (size_t)n * capacity * 9;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int h = 0; h < 3; ++h) {
        const int start = carry_s;
        __syncthreads();
        for (int b = 0; b < c[h]; b += kCollectThreads) {
            const int i = b + tid;
            bool sel = false;
            if (i < c[h]) {
                if (bug_compat) {
                    for (int k = 0; k < 3; ++k) sel = sel || (i < c[k] && kp[off[k] + i] != 0);
                } else {
                    sel = kp[off[h] + i] != 0;
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, sel);
            if (lane == 0) wsum_s[warp] = __popc(bal);
            __syncthreads();
            int base = carry_s;
            for (int w = 0; w < warp; ++w) base += wsum_s[w];
            if (sel) {
                const int pos = base + __popc(bal & ((1u << lane) - 1u));
                const float4 r0 = *reinterpret_cast<const float4*>(rows + (size_t)(off[h] + i) * YV3_CAND_COLS);
                const float4 r1 = *reinterpret_cast<const float4*>(rows + (size_t)(off[h] + i) * YV3_CAND_COLS + 4);
                double* o = out + (size_t)pos * 9;
                o[0] = __dmul_rn((double)r0.x, tw); o[1] = __dmul_rn((double)r0.y, th);
                o[2] = __dmul_rn((double)r0.z, tw); o[3] = __dmul_rn((double)r0.w, th);
                o[4] = (double)r1.x; o[5] = (double)r1.y; o[6] = (double)r1.z; o[7] = (double)r1.w;
                o[8] = (double)i;
            }
            __syncthreads();
            if (tid == 0) {
                int tot = 0;
                for (int w = 0; w < kCollectThreads / 32; ++w) tot += wsum_s[w];
                carry_s += tot;
            }
            __syncthreads();
        }
        if (tid == 0) counts_out[n * 3 + h] = carry_s - start;
        __syncthreads();
    }
}

cudaError_t launch_collect(const float* cand, const int32_t* counts, const unsigned char* keep, int n, int capacity,
                           double tw, double th, int bug_compat, double* rows_out, int32_t* counts_out, cudaStream_t stream) {
    collect_kernel<<<n, kCollectThreads, 0, stream>>>(cand, counts, keep, capacity, tw, th, bug_compat, rows_out, counts_out);
    return cudaGetLastError();
}

// ---- launchers ------------------------------------------------------------------------------------------
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) sticks to the function on a device: it leaves the launch path once a
// launch has asked for at least as much (a few microseconds of host time per call otherwise, on every step)
static cudaError_t set_smem_once(const void* func, size_t bytes) {
    struct Seen { const void* f; int dev; size_t bytes; };
    static Seen seen[64];
    static int nseen = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    for (int i = 0; i < nseen; ++i)
        if (seen[i].f == func && seen[i].dev == dev) {
            if (seen[i].bytes >= bytes) return cudaSuccess;
            cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
            if (e == cudaSuccess) seen[i].bytes = bytes;
            return e;
        }
    cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    // a kernel launched from the device takes its L1 / shared-memory split from the function's preference alone
    if (e == cudaSuccess) (void)cudaFuncSetAttribute(func, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e == cudaSuccess && nseen < 64) seen[nseen++] = Seen{func, dev, bytes};
    return e;
}

static ScoreTiles score_tiles(const Geo& g, const float* predicts, int n, float thresh, int tr_force = 0) {
    ScoreTiles q;
    q.total_rec = (long long)n * g.rpi;
    long long tr = 32;
    tr = min(tr, (long long)(200 * 1024) / ((long long)g.L * 4));   // a very long record: fewer records per warp
    tr = min(tr, (long long)g.rpi);                               // a tile touches at most two images
    tr &= ~3LL;
    q.tr = tr_force ? tr_force : (int)max(tr, 0LL);
    q.tile_floats = (int)(align_up((size_t)max(q.tr, 1) * g.L * 4, 128) / 4);
    // one warp (one tile) per CTA: the tile buffers fill the SM's shared memory either way (~19 per SM at L = 85), but
    // a finished warp's buffer goes back to the block scheduler at once instead of waiting for three siblings
    // (measured at cfg2 on B200: 52.7 us with four warps per CTA, 46.6 us with one; profiles/r01_score_variants.txt)
    q.wpc = 1;
    if (const char* e = getenv("YV3_SCORE_WPC")) {                 // developer override
        const long long fit = max(1LL, min((long long)(kScoreThreads / 32), (long long)45056 / ((long long)q.tile_floats * 4)));
        const int v = atoi(e);
        if (v >= 1 && v <= fit) q.wpc = v;
    }
    q.num_tiles = q.tr ? (int)((q.total_rec + q.tr - 1) / q.tr) : 0;
    q.tma = (((uintptr_t)predicts & 15u) == 0) ? 1 : 0;
    q.early = (thresh > 2.0f * g.lo) ? 1 : 0;
    return q;
}

ScoreTiles packed_score_tiles(const Geo& g, const float* predicts, int n, float thresh) {
    return score_tiles(g, predicts, n, thresh);
}

// records per tile when every head tensor is swept on its own: the smallest of the three (a tile touches two images at most)
int heads_tr(const Geo& g, int n) {
    int tr = 32;
    for (int h = 0; h < 3; ++h) tr = min(tr, score_tiles(head_geo(g, h), nullptr, n, 0.f).tr);
    return tr;
}

// tiles the workspace must hold: the packed sweep or the three per-head sweeps, whichever is more
int score_num_tiles(const Geo& g, int n) {
    const int packed = score_tiles(g, nullptr, n, 0.f).num_tiles;
    const int tr = heads_tr(g, n);
    int heads = 0;
    if (tr >= 4) for (int h = 0; h < 3; ++h) heads += score_tiles(head_geo(g, h), nullptr, n, 0.f, tr).num_tiles;
    return max(packed, heads) + 4;
}

// pred[0] alone: the packed tensor; unpacked: one tensor per head (N, H, W, B*L), each swept under its single-head geometry
cudaError_t launch_filter_packed(const Geo& g, const float* const pred[3], bool unpacked, int n, float thresh, float* cand,
                                 int capacity, int32_t* counts, int32_t* status, StageRow* stage, int32_t* tcount,
                                 bool sweep, bool gather, cudaStream_t stream) {
    GatherSrc src;
    memset(&src, 0, sizeof(src));
    const int tr = unpacked ? heads_tr(g, n) : score_tiles(g, pred[0], n, thresh).tr;
    if (tr < 4) return cudaErrorInvalidValue;
    int tbase = 0, gtiles = 0;
    for (int h = 0; h < (unpacked ? 3 : 1); ++h) {
        const Geo gh = unpacked ? head_geo(g, h) : g;
        const ScoreTiles q = score_tiles(gh, pred[h], n, thresh, tr);
        const size_t smem = (size_t)q.wpc * q.tile_floats * 4;
        if (smem > 220 * 1024) return cudaErrorInvalidValue;
        cudaError_t err = set_smem_once((const void*)score_tile_kernel, smem);
        if (err != cudaSuccess) return err;
        const int grid = (q.num_tiles + q.wpc - 1) / q.wpc;
        if (sweep) {
            score_tile_kernel<<<grid, 32 * q.wpc, smem, stream>>>(gh, pred[h], q, thresh, stage + (size_t)tbase * tr, tcount + (size_t)tbase * 8);
            err = cudaGetLastError();
            if (err != cudaSuccess) return err;
        }
        for (int k = (unpacked ? h : 0); k < (unpacked ? h + 1 : 3); ++k) {
            src.pred[k] = pred[h];
            src.rpi[k] = gh.rpi; src.rpr[k] = gh.rpr; src.seg[k] = unpacked ? 0 : g.seg_base[k];
            src.tbase[k] = tbase;
            gtiles += (int)((gh.rpi + tr - 1) / tr) + 2;
        }
        tbase += q.num_tiles;
    }
    if (!gather) return cudaSuccess;
    const size_t gsmem = (size_t)3 * gtiles * sizeof(int);
    if (gsmem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t err = set_smem_once((const void*)gather_kernel, gsmem);
    if (err != cudaSuccess) return err;
    gather_kernel<<<n, kGatherThreads, gsmem, stream>>>(g, src, tr, stage, tcount, cand, capacity, counts, status);
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
    cudaError_t err = cudaSuccess;
    // keep is NOT cleared here: every row below the list's count is written (0 or 1) by exactly one of the kernels below,
    // and rows beyond the count are never read (yv3_collect_kept and every caller stop at the counts)
    // few classes (the single-class heads): lists of thousands of ONE class are the rule, the general path sorts them
    // whole, and a 16 384-key bitonic sort in global memory took 180 us per step: those launches get the large reservation
    const int n2s = min(cap2, num_classes <= 8 ? 16384 : kSortSmemKeys);
    const size_t smem = (size_t)n2s * 8 + (size_t)kBucketClasses * 4 + (size_t)n2s * 2;
    const bool two_per_sm = n > 148 && smem <= 100 * 1024 && !getenv("YV3_SORT_ONE_PER_SM");
    err = two_per_sm ? set_smem_once((const void*)sort_kernel<2>, smem) : set_smem_once((const void*)sort_kernel<1>, smem);
    if (err != cudaSuccess) return err;
    err = cudaMemsetAsync(ws.wl_count, 0, 3 * sizeof(int32_t), stream);          // two queue counters, the dispatch ticket
    if (err != cudaSuccess) return err;
    int2* wl_large = ws.wl + (size_t)n * capacity;
    static int skip = -1;                                     // developer switch (timing experiments only): YV3_NMS_SKIP bits
    if (skip < 0) { const char* e = getenv("YV3_NMS_SKIP"); skip = e ? atoi(e) : 0; }
    // many classes: the queue-driven kernels are launched from the device, and only when something is queued
    static int dev_warps = -1, host_ctas = -1;                // developer switches (timing experiments only)
    if (dev_warps < 0) {
        const char* e = getenv("YV3_WARP_NMS_DEV_WARPS");
        dev_warps = e ? max(1, min(atoi(e), kWarpNmsDevThreads / 32)) : kWarpNmsDevThreads / 32;
        e = getenv("YV3_WARP_NMS_HOST_CTAS");
        host_ctas = e ? max(1, atoi(e)) : 24;
    }
    const int dispatch = (num_classes > 8 && !(skip & 8)) ? dev_warps : 0;
    err = set_smem_once((const void*)nms_kernel<kNmsLarge>, nms_smem_bytes(kNmsLarge));
    if (err != cudaSuccess) return err;
    const size_t wsmem = sizeof(WarpNmsSmem) * kWarpNmsWarps;
    err = set_smem_once((const void*)nms_warp_kernel, sizeof(WarpNmsSmem) * (kWarpNmsDevThreads / 32));
    if (err != cudaSuccess) return err;
    if (!(skip & 1))
    if (two_per_sm) sort_kernel<2><<<n, kSortThreads, smem, stream>>>(cand, count, count_cols, capacity, cap2, n2s, thresh, ws.keys, ws.sbox,
                                                   ws.sidx, keep, ws.wl, wl_large, ws.wl_count, ws.kspill, dispatch, num_classes > 8 ? 1 : 0);
    else sort_kernel<1><<<n, kSortThreads, smem, stream>>>(cand, count, count_cols, capacity, cap2, n2s, thresh, ws.keys, ws.sbox,
                                                   ws.sidx, keep, ws.wl, wl_large, ws.wl_count, ws.kspill, dispatch, num_classes > 8 ? 1 : 0);
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    if (dispatch) return cudaSuccess;
    if (!(skip & 2))
    nms_warp_kernel<<<148 * host_ctas, kWarpNmsThreads, wsmem, stream>>>(ws.keys, count, count_cols, capacity, cap2, thresh, cand,
                                                                 keep, ws.wl, ws.wl_count);
    err = cudaGetLastError();
    if (err != cudaSuccess) return err;
    static int clusters = -1;
    if (clusters < 0) {
        clusters = 0;
        if (cudaFuncSetAttribute(nms_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)nms_cluster_smem_bytes()) == cudaSuccess) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(kClu);
            q.blockDim = dim3(kCluT);
            q.dynamicSmemBytes = nms_cluster_smem_bytes();
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = kClu; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            q.attrs = at;
            q.numAttrs = 1;
            int nc = 0;
            if (cudaOccupancyMaxActiveClusters(&nc, nms_cluster_kernel, &q) == cudaSuccess) clusters = nc;
        }
        (void)cudaGetLastError();
        if (getenv("YV3_NO_CLUSTER")) clusters = 0;                  // developer switch
    }
    // Segments longer than 256 candidates need an image with more than 256 candidates of ONE class: the rule with a
    // single-class head and a low threshold, a rarity with tens of classes.  The cluster kernel's CTAs (1024 threads,
    // 147 KB, six at a time on one GPC) cost the SMs they land on a dozen of the sweep's CTAs even when the queue is
    // empty -- launched on every step that was most of the gap between the step and its sweep -- so with many classes the
    // queue is walked by four CTAs of the single-CTA kernel instead: correct for any input, slow only for the rare one.
    if (skip & 4) return cudaSuccess;
    if (num_classes > 8) {
        nms_kernel<kNmsLarge><<<4, kNmsLarge, nms_smem_bytes(kNmsLarge), stream>>>(
            ws.keys, count, count_cols, capacity, cap2, thresh, ws.sbox, ws.sidx, ws.kspill, keep, wl_large, ws.wl_count + 1);
        return cudaGetLastError();
    }
    if (clusters >= 1) {
        cudaLaunchConfig_t q = {};
        q.gridDim = dim3(kClu * min(clusters, 32));
        q.blockDim = dim3(kCluT);
        q.dynamicSmemBytes = nms_cluster_smem_bytes();
        q.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = kClu; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        q.attrs = at;
        q.numAttrs = 1;
        const unsigned long long* a0 = ws.keys;
        const int2* a10 = wl_large;
        const int32_t* a11 = ws.wl_count + 1;
        return cudaLaunchKernelEx(&q, nms_cluster_kernel, a0, count, count_cols, capacity, cap2, thresh,
                                  (const float4*)ws.sbox, (const int32_t*)ws.sidx, ws.kspill, keep, a10, a11);
    }
    nms_kernel<kNmsLarge><<<148, kNmsLarge, nms_smem_bytes(kNmsLarge), stream>>>(
        ws.keys, count, count_cols, capacity, cap2, thresh, ws.sbox, ws.sidx, ws.kspill, keep, wl_large, ws.wl_count + 1);
    return cudaGetLastError();
}

}  // namespace yv3
