// mAP accumulation after NMS, on the device.
//
// The reference stops at a TODO: run.py:78-79 / README.md:137 point to the third-party tool Cartucho/mAP (not vendored,
// no version pinned) for "AP, mAP from the predictions".  This file implements that tool's published algorithm
// (main.py: VOC2012 all-point AP) on the rows yv3_collect_kept leaves on the device, so that evaluation needs no
// per-box host code:
//   match     per class, detections by confidence descending (stable: file order, i.e. image order then row order);
//             a detection's candidate is the ground-truth box of its image and class with the LARGEST IoU (first one
//             on ties), IoU on pixel boxes with the tool's inclusive "+1" widths; it is a true positive iff that IoU
//             >= min_overlap (0.5) and no earlier detection has claimed the same ground-truth box, else a false positive;
//   AP        cumulative tp / fp -> recall = tp / #gt, precision = tp / (tp + fp); precision made monotone from the
//             right; AP = sum over the points where recall changes of (recall step) * precision;  mAP = mean over the
//             classes that have ground truth.
// Because a detection's candidate box does not depend on which boxes are already claimed, "no earlier detection has
// claimed it" is "this detection ranks first among the detections whose candidate is that box and whose IoU passes":
// an atomicMax per ground-truth box on (score, reversed row) -- images are independent, one CTA each.
// All arithmetic that the tool does in Python floats is done in fp64, in the tool's operation order.
#include "yv3_kernels.cuh"

namespace yv3 {

struct MapRecord {
    float score;
    int32_t cls_tp;              // class << 1 | tp
    unsigned long long order;    // image << 24 | row: the tool's file order, for stable ties
};
static_assert(sizeof(MapRecord) == 16, "MapRecord must stay 16 bytes");

struct MapState {
    int32_t count;               // records appended so far
    int32_t max_det, num_classes, overflow;
    // then int32 gt_count[num_classes], int32 cls_count[num_classes], int32 cls_off[num_classes + 1], pad to 256,
    // MapRecord rec[max_det], MapRecord sorted[max_det] (after finalize: per record (precision, recall) doubles)
};
__host__ __device__ inline size_t map_ints_off() { return 16; }
__host__ __device__ inline size_t map_rec_off(int num_classes) { return align_up(16 + sizeof(int32_t) * (3 * (size_t)num_classes + 1), 256); }
__host__ __device__ inline size_t map_state_bytes(int max_det, int num_classes) {
    return map_rec_off(num_classes) + 2 * sizeof(MapRecord) * (size_t)max_det;
}

__device__ __forceinline__ unsigned ordered_u32(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int kMapThreads = 128;
constexpr int kMapMaxGt = YV3_MAX_GT;

// IoU of the tool: inclusive pixel widths (+1) when `inc` is 1.0, plain widths when 0.0; -1 when the boxes do not overlap
__device__ __forceinline__ double map_iou(const double* bb, const double4& gt, double inc) {
    const double bx0 = fmax(bb[0], gt.x), by0 = fmax(bb[1], gt.y), bx1 = fmin(bb[2], gt.z), by1 = fmin(bb[3], gt.w);
    const double iw = __dadd_rn(__dsub_rn(bx1, bx0), inc), ih = __dadd_rn(__dsub_rn(by1, by0), inc);
    if (!(iw > 0.0 && ih > 0.0)) return -1.0;
    const double a = __dmul_rn(__dadd_rn(__dsub_rn(bb[2], bb[0]), inc), __dadd_rn(__dsub_rn(bb[3], bb[1]), inc));
    const double b = __dmul_rn(__dadd_rn(__dsub_rn(gt.z, gt.x), inc), __dadd_rn(__dsub_rn(gt.w, gt.y), inc));
    const double inter = __dmul_rn(iw, ih);
    const double ua = __dsub_rn(__dadd_rn(a, b), inter);
    return __ddiv_rn(inter, ua);
}

__global__ void __launch_bounds__(kMapThreads) map_match_kernel(const double* __restrict__ rows, const int32_t* __restrict__ kcounts,
                                                                int capacity, const float* __restrict__ targets, int m,
                                                                double img_w, double img_h, int image_base, double min_overlap,
                                                                double inc, int score_col, int class_agnostic,
                                                                unsigned char* __restrict__ state) {
    __shared__ double4 gt_s[kMapMaxGt];
    __shared__ int gcls_s[kMapMaxGt];
    __shared__ unsigned long long claim_s[kMapMaxGt];
    __shared__ int base_s;
    MapState* st = reinterpret_cast<MapState*>(state);
    int32_t* gt_count = reinterpret_cast<int32_t*>(state + map_ints_off());
    MapRecord* rec = reinterpret_cast<MapRecord*>(state + map_rec_off(st->num_classes));
    const int n = blockIdx.x, tid = threadIdx.x;
    const int K = min(capacity, max(0, kcounts[n * 3] + kcounts[n * 3 + 1] + kcounts[n * 3 + 2]));
    const double* r0 = rows + (size_t)n * capacity * 9;
    for (int g = tid; g < m; g += kMapThreads) {
        const float* t = targets + ((size_t)n * m + g) * 5;
        const double x = (double)t[0], y = (double)t[1], w = (double)t[2], h = (double)t[3];
        int c = -1;
        if (t[0] >= 0.0f) {                                           // padding rows are -1 (dataset/file_util.py:97)
            c = class_agnostic ? 0 : (int)t[4];
            if (c >= 0 && c < st->num_classes) atomicAdd(&gt_count[c], 1); else c = -1;
        }
        gcls_s[g] = c;
        const double hw = __dmul_rn(w, 0.5), hh = __dmul_rn(h, 0.5);
        gt_s[g] = make_double4(__dmul_rn(__dsub_rn(x, hw), img_w), __dmul_rn(__dsub_rn(y, hh), img_h),
                               __dmul_rn(__dadd_rn(x, hw), img_w), __dmul_rn(__dadd_rn(y, hh), img_h));
        claim_s[g] = 0ULL;
    }
    if (tid == 0) base_s = K > 0 ? atomicAdd(&st->count, K) : 0;
    __syncthreads();
    // pass 1: every detection finds its candidate box; passing ones bid for it
    for (int pass = 0; pass < 2; ++pass) {
        for (int d = tid; d < K; d += kMapThreads) {
            const double* bb = r0 + (size_t)d * 9;
            const int c = class_agnostic ? 0 : (int)bb[6];
            double ovmax = -1.0;
            int best = -1;
            for (int g = 0; g < m; ++g) {
                if (gcls_s[g] != c) continue;
                const double ov = map_iou(bb, gt_s[g], inc);
                if (ov > ovmax) { ovmax = ov; best = g; }
            }
            const float score = (float)bb[score_col];
            const unsigned long long bid = ((unsigned long long)ordered_u32(score) << 32) | (unsigned long long)(0xffffffffu - (unsigned)d);
            const bool passes = best >= 0 && ovmax >= min_overlap;
            if (pass == 0) {
                if (passes) atomicMax(&claim_s[best], bid);
            } else {
                const int tp = (passes && claim_s[best] == bid) ? 1 : 0;
                const int pos = base_s + d;
                if (pos < st->max_det) {
                    MapRecord o;
                    o.score = score;
                    o.cls_tp = (c << 1) | tp;
                    o.order = ((unsigned long long)(unsigned)(image_base + n) << 24) | (unsigned long long)d;
                    rec[pos] = (c >= 0 && c < st->num_classes) ? o : MapRecord{0.f, -2, 0ULL};
                } else {
                    st->overflow = 1;
                }
            }
        }
        __syncthreads();
    }
}

// ---- finalize ---------------------------------------------------------------------------------------------------------
__global__ void map_bucket_kernel(unsigned char* __restrict__ state, int phase) {
    MapState* st = reinterpret_cast<MapState*>(state);
    const int C = st->num_classes, total = min(st->count, st->max_det);
    int32_t* cls_count = reinterpret_cast<int32_t*>(state + map_ints_off()) + C;
    int32_t* cls_off = cls_count + C;
    MapRecord* rec = reinterpret_cast<MapRecord*>(state + map_rec_off(C));
    MapRecord* sorted = rec + st->max_det;
    const int gid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    if (phase == 0) {
        for (int i = gid; i < C; i += stride) cls_count[i] = 0;
    } else if (phase == 1) {
        for (int i = gid; i < total; i += stride) {
            const int c = rec[i].cls_tp >> 1;
            if (c >= 0) atomicAdd(&cls_count[c], 1);
        }
    } else if (phase == 2) {
        if (gid == 0) {
            int acc = 0;
            for (int c = 0; c < C; ++c) { cls_off[c] = acc; acc += cls_count[c]; cls_count[c] = 0; }
            cls_off[C] = acc;
        }
    } else {
        for (int i = gid; i < total; i += stride) {
            const int c = rec[i].cls_tp >> 1;
            if (c >= 0) sorted[cls_off[c] + atomicAdd(&cls_count[c], 1)] = rec[i];
        }
    }
}

// key order: score descending, then the tool's file order ascending
__device__ __forceinline__ bool map_before(const MapRecord& a, const MapRecord& b) {
    if (a.score != b.score) return a.score > b.score;
    return a.order < b.order;
}

constexpr int kMapSortThreads = 1024;
// One CTA per class: bitonic sort of the class's records in global memory, then the tool's AP by one thread in its
// operation order (precision / recall arrays go into the spent unsorted buffer).
__global__ void __launch_bounds__(kMapSortThreads) map_ap_kernel(unsigned char* __restrict__ state, double* __restrict__ ap_out,
                                                                 int32_t* __restrict__ gt_out) {
    MapState* st = reinterpret_cast<MapState*>(state);
    const int C = st->num_classes, c = blockIdx.x, tid = threadIdx.x;
    const int32_t* gt_count = reinterpret_cast<const int32_t*>(state + map_ints_off());
    const int32_t* cls_off = gt_count + 2 * C;
    MapRecord* rec = reinterpret_cast<MapRecord*>(state + map_rec_off(C));
    MapRecord* seg = rec + st->max_det + cls_off[c];
    double2* pr = reinterpret_cast<double2*>(rec) + cls_off[c];                  // (precision, recall) per record
    const int n = cls_off[c + 1] - cls_off[c];
    int n2 = 1;
    while (n2 < n) n2 <<= 1;
    // bitonic network in its all-ascending form (a mirrored "flip" step, then half-cleaners): every comparator puts the
    // earlier record at the lower index, so the virtual padding beyond n (sorts last) never has to move
    for (int k = 2; k <= n2; k <<= 1) {
        for (int i = tid; i < (n2 >> 1); i += kMapSortThreads) {
            const int half = k >> 1, blk = i / half, off = i - blk * half;
            const int lo = blk * k + off, hi = blk * k + k - 1 - off;
            if (hi < n) {
                const MapRecord a = seg[lo], b = seg[hi];
                if (map_before(b, a)) { seg[lo] = b; seg[hi] = a; }
            }
        }
        __syncthreads();
        for (int j = k >> 2; j > 0; j >>= 1) {
            for (int i = tid; i < (n2 >> 1); i += kMapSortThreads) {
                const int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1)), hi = lo | j;
                if (hi < n) {
                    const MapRecord a = seg[lo], b = seg[hi];
                    if (map_before(b, a)) { seg[lo] = b; seg[hi] = a; }
                }
            }
            __syncthreads();
        }
    }
    if (tid != 0) return;
    const int gtc = gt_count[c];
    if (gt_out) gt_out[c] = gtc;
    double tp = 0.0, fp = 0.0;
    for (int i = 0; i < n; ++i) {
        if (seg[i].cls_tp & 1) tp += 1.0; else fp += 1.0;
        pr[i] = make_double2(__ddiv_rn(tp, __dadd_rn(fp, tp)), gtc > 0 ? __ddiv_rn(tp, (double)gtc) : 0.0);
    }
    double run = 0.0;                                               // mpre[n + 1] = 0
    for (int i = n - 1; i >= 0; --i) {
        run = fmax(pr[i].x, run);
        pr[i].x = run;
    }
    double ap = 0.0, prev = 0.0;                                    // mrec[0] = 0
    for (int i = 0; i < n; ++i) {
        const double r = pr[i].y;
        if (r != prev) ap = __dadd_rn(ap, __dmul_rn(__dsub_rn(r, prev), pr[i].x));
        prev = r;
    }
    ap_out[c] = gtc > 0 ? ap : -1.0;                                // -1: the class has no ground truth (not in the mean)
}

__global__ void map_mean_kernel(const unsigned char* __restrict__ state, const double* __restrict__ ap, double* __restrict__ map_out) {
    const MapState* st = reinterpret_cast<const MapState*>(state);
    double sum = 0.0;
    int cnt = 0;
    for (int c = 0; c < st->num_classes; ++c)
        if (ap[c] >= 0.0) { sum = __dadd_rn(sum, ap[c]); ++cnt; }
    map_out[0] = cnt ? __ddiv_rn(sum, (double)cnt) : 0.0;
    map_out[1] = (double)st->overflow;
}

}  // namespace yv3

using namespace yv3;

extern "C" {

int yv3_map_state_bytes(int max_detections, int num_classes, size_t* bytes) {
    if (!bytes) return YV3_ERR_NULL;
    if (max_detections < 1 || num_classes < 1 || num_classes > (1 << 20)) return YV3_ERR_SHAPE;
    *bytes = map_state_bytes(max_detections, num_classes);
    return YV3_OK;
}

int yv3_map_reset(void* state, size_t state_bytes, int max_detections, int num_classes, void* stream) {
    if (!state) return YV3_ERR_NULL;
    if (max_detections < 1 || num_classes < 1) return YV3_ERR_SHAPE;
    if (state_bytes < map_state_bytes(max_detections, num_classes)) return YV3_ERR_WORKSPACE;
    if (((uintptr_t)state & 15u) != 0) return YV3_ERR_ALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(state, 0, map_rec_off(num_classes), st) != cudaSuccess) return YV3_ERR_CUDA;
    const int32_t head[4] = {0, max_detections, num_classes, 0};
    if (cudaMemcpyAsync(state, head, sizeof(head), cudaMemcpyHostToDevice, st) != cudaSuccess) return YV3_ERR_CUDA;
    return YV3_OK;
}

int yv3_map_accumulate(const double* kept_rows, const int32_t* kept_counts, int n, int32_t capacity, const float* targets,
                       int m, double image_w, double image_h, int image_base, double min_overlap, int pixel_inclusive,
                       int score_col, int class_agnostic, void* state, void* stream) {
    if (!kept_rows || !kept_counts || !targets || !state) return YV3_ERR_NULL;
    if (n < 1 || capacity < 1 || m < 1 || m > kMapMaxGt || (score_col != 4 && score_col != 7)) return YV3_ERR_SHAPE;
    if (image_base < 0 || (long long)image_base + n >= (1LL << 31) || capacity > (1 << 24)) return YV3_ERR_CAPACITY;
    map_match_kernel<<<n, kMapThreads, 0, (cudaStream_t)stream>>>(kept_rows, kept_counts, capacity, targets, m, image_w, image_h,
                                                                 image_base, min_overlap, pixel_inclusive ? 1.0 : 0.0, score_col,
                                                                 class_agnostic ? 1 : 0, (unsigned char*)state);
    return cudaGetLastError() == cudaSuccess ? YV3_OK : YV3_ERR_CUDA;
}

int yv3_map_finalize(void* state, int num_classes, double* ap_out, int32_t* gt_count_out, double* map_out2, void* stream) {
    if (!state || !ap_out || !map_out2) return YV3_ERR_NULL;
    if (num_classes < 1) return YV3_ERR_SHAPE;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* s = (unsigned char*)state;
    map_bucket_kernel<<<64, 256, 0, st>>>(s, 0);
    map_bucket_kernel<<<296, 256, 0, st>>>(s, 1);
    map_bucket_kernel<<<1, 32, 0, st>>>(s, 2);
    map_bucket_kernel<<<296, 256, 0, st>>>(s, 3);
    map_ap_kernel<<<num_classes, kMapSortThreads, 0, st>>>(s, ap_out, gt_count_out);
    map_mean_kernel<<<1, 1, 0, st>>>(s, ap_out, map_out2);
    return cudaGetLastError() == cudaSuccess ? YV3_OK : YV3_ERR_CUDA;
}

}  // extern "C"
