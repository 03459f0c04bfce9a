// The fused, memory-bound pass of the head loss: decode + max-IoU + masks + noobj / rectified / class
// loss terms + the complete gradient tensor, one sweep over the packed predicts.
//
// Follows yolov3/yolov3_loss.py: the all-anchor part of _calc_iou :267-294, the masks of
// _single_head_loss :328-332, the noobj term :335-338, the class term :361-364, the rectified term
// :140-164, the batch mean :112 and the detail rows :115-138.  The backward is the closed form TF
// autodiff yields for those operators; no gradient flows through IoU or masks.
//
// Shape of the kernel (B200), v2 -- profiles/r01_v1_ncu_summary.md says why v1 (two 87 KB tile images per
// SM pushed out by TMA, one 8-warp CTA per SM) was latency-bound:
//  * the packed tensor is a flat array of records of L floats; a tile = TR (<= 256) consecutive records,
//    whose byte offset is a multiple of 16, so the gradient of a tile is one 16-B aligned span;
//  * nothing of the gradient is staged: a tile's span is first covered with 128-bit streaming ZERO stores
//    (fire-and-forget, issued before the arithmetic so they overlap it), and after the CTA barrier every
//    thread overwrites the five box/conf floats of its own record with their values.  L-5 of a record's L
//    gradient floats are zeros, so this is ~94 % of the bytes at ~3 instructions per 16 B;
//  * a thread owns one record: it reads the record's 5 box/conf logits (class logits of non-responsible
//    anchors are never read, their gradient is identically 0), decodes, takes the max IoU over the
//    image's labels held in shared memory, accumulates the loss terms and stores 5 gradient values;
//  * shared memory holds only the decoded labels of the <= 2 images a tile touches and the list of owned
//    entries falling into the tile, double-buffered by tile parity => ONE barrier per tile and several
//    CTAs per SM, so the 20-B logit gathers of one CTA overlap the arithmetic and stores of the others;
//  * the few records that own a label get the owner terms added in label order by their thread, and
//    their softmax-CE class gradient added on top of the zeros by a warp;
//  * loss sums: registers -> warp shuffles -> one partial row per CTA -> the last CTA to finish adds
//    the rows in index order (double) and writes the 18 detail values + their sum.  No float atomics.
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kSlots = 2;             // images a tile may touch (TR <= records per image)

struct DenseBuf {                     // one per tile parity
    float* gt[kSlots];                // [3][m][5]  (x0, y0, x1, y1, area) of the valid labels, label order
    int* list_rec;                    // [cap] tile-relative record of an owned entry
    int* list_idx;                    // [cap] index into own[]
    int* meta;                        // [0..1] valid-label count per slot, [2] entries in the list
};
struct DenseSmem {
    DenseBuf buf[2];
    float* red;                       // [kWarps][12]
};

__host__ __device__ inline size_t dense_smem_bytes(int m) {
    const size_t cap = (size_t)kSlots * 3 * m;
    size_t per = kSlots * align_up((size_t)3 * m * 5 * 4, 16) + 2 * align_up(cap * 4, 16) + 16;
    return 2 * per + kWarps * 12 * 4 + 16;
}

__device__ __forceinline__ DenseSmem carve(unsigned char* base, int m) {
    DenseSmem s;
    size_t o = (16 - ((size_t)smem_u32(base) & 15)) & 15;
    auto take = [&](size_t bytes) {
        unsigned char* p = base + o;
        o += (bytes + 15) / 16 * 16;
        return p;
    };
    const size_t cap = (size_t)kSlots * 3 * m;
#pragma unroll
    for (int b = 0; b < 2; ++b) {
        s.buf[b].gt[0] = (float*)take((size_t)3 * m * 5 * 4);
        s.buf[b].gt[1] = (float*)take((size_t)3 * m * 5 * 4);
        s.buf[b].list_rec = (int*)take(cap * 4);
        s.buf[b].list_idx = (int*)take(cap * 4);
        s.buf[b].meta = (int*)take(16);
    }
    s.red = (float*)take(kWarps * 12 * 4);
    return s;
}

__global__ void __launch_bounds__(kThreads, 3) loss_dense_kernel(const __grid_constant__ Geo g, const DenseArgs a) {
    extern __shared__ unsigned char smem_raw[];
    const bool with_grad = a.grad != nullptr && !(a.debug & 8);
    const DenseSmem s = carve(smem_raw, a.m);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int L = g.L, C = g.C, m = a.m, tr = a.tr;
    const int cap_entries = 3 * m;

    float acc_noobj[3] = {0.f, 0.f, 0.f}, acc_rect[3] = {0.f, 0.f, 0.f}, acc_cls[3] = {0.f, 0.f, 0.f};
    const float thr = g.iou_thresh;
    const float nf = (float)a.n_global;
    const float one = 1.0f;

    float cur[5];
    auto load_rec = [&](int tile, float (&dst)[5]) {
        const long long grec = (long long)tile * tr + tid;
        const bool ok = tile < a.num_tiles && tid < tr && grec < a.total_rec;
        const float* p = a.predicts + (size_t)(ok ? grec : 0) * L;
#pragma unroll
        for (int c = 0; c < 5; ++c) dst[c] = ok ? __ldg(p + c) : 0.f;
    };
    load_rec(blockIdx.x, cur);

    int parity = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, parity ^= 1) {
        const DenseBuf& B = s.buf[parity];
        const long long rec0 = (long long)tile * tr;
        const int nrec = (int)min((long long)tr, a.total_rec - rec0);
        const int n_first = (int)(rec0 / g.rpi);
        const int n_last = (int)((rec0 + nrec - 1) / g.rpi);
        const int nslots = n_last - n_first + 1;          // 1 or 2
        const int nfl = nrec * L;

        // ---- zeros first: the whole span, 16 B per store, nothing waits for them ----------------------
        if (with_grad) {
            float4* gdst = reinterpret_cast<float4*>(a.grad + (size_t)rec0 * L);
            const int nchunks = nfl >> 2;
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
            for (int c = tid; c < nchunks; c += kThreads) __stcs(gdst + c, z4);
            // only the very last tile of the tensor can end off a 16-byte boundary
            if (tid < (nfl & 3)) a.grad[(size_t)rec0 * L + (nchunks << 2) + tid] = 0.f;
        }

        // ---- tile setup (into this parity's buffers) ---------------------------------------------------
        // warp sl (< nslots): valid labels of image slot sl -> per-head boxes, order kept (:239-247);
        // warp 2: the owned entries that fall into this tile, order kept
        if (warp < nslots) {
            const int sl = warp;
            float* gts = B.gt[sl];
            const float* lab = a.targets + (size_t)(n_first + sl) * m * 5;
            int count = 0;
            for (int base = 0; base < m; base += 32) {
                const int j = base + lane;
                float xn = -1.f, yn = 0.f, wn = 0.f, hn = 0.f;
                if (j < m) {
                    const float* t = lab + j * 5;
                    xn = __ldg(t); yn = __ldg(t + 1); wn = __ldg(t + 2); hn = __ldg(t + 3);
                }
                const bool valid = (j < m) && (__fmul_rn(xn, g.fW[0]) >= 0.0f);
                const unsigned mask = __ballot_sync(0xffffffffu, valid);
                const int pos = count + __popc(mask & ((1u << lane) - 1u));
                if (valid) {
#pragma unroll
                    for (int h = 0; h < 3; ++h) {
                        const GtDec T = decode_label(g, h, xn, yn, wn, hn);
                        float* o = gts + ((size_t)h * m + pos) * 5;
                        o[0] = T.box.x0; o[1] = T.box.y0; o[2] = T.box.x1; o[3] = T.box.y1; o[4] = T.box.area;
                    }
                }
                count += __popc(mask);
            }
            if (lane == 0) B.meta[sl] = count;
        } else if (warp == 2) {
            int count = 0;
            if (!(a.debug & 4)) {
                for (int sl = 0; sl < nslots; ++sl) {
                    const int n_img = n_first + sl;
                    const long long img_rec0 = (long long)n_img * g.rpi;
                    for (int base = 0; base < cap_entries; base += 32) {
                        const int e = base + lane;
                        int rec = -1;
                        if (e < cap_entries) rec = a.own[(size_t)n_img * cap_entries + e].rec;
                        const long long grec = img_rec0 + rec;
                        const bool in = rec >= 0 && grec >= rec0 && grec < rec0 + nrec;
                        const unsigned mask = __ballot_sync(0xffffffffu, in);
                        const int pos = count + __popc(mask & ((1u << lane) - 1u));
                        if (in) {
                            B.list_rec[pos] = (int)(grec - rec0);
                            B.list_idx[pos] = n_img * cap_entries + e;
                        }
                        count += __popc(mask);
                    }
                }
            }
            if (lane == 0) B.meta[2] = count;
        }
        // prefetch the next tile's logits while this one is processed
        float nxt[5];
        load_rec(tile + gridDim.x, nxt);
        __syncthreads();          // the tile's one barrier: buffers ready, zero stores ordered before value stores
        const int nl = B.meta[2];

        // ---- record phase: thread = record ---------------------------------------------------------
        if (tid < nrec) {
            const long long grec = rec0 + tid;
            const int n_img = (int)(grec / g.rpi);
            const int r_img = (int)(grec - (long long)n_img * g.rpi);
            const RecLoc p = locate(g, r_img);
            const BoxDec d = decode_box(g, p, cur[0], cur[1], cur[2], cur[3], cur[4]);
            const float parea = __fmul_rn(d.w, d.h);                                       // :267
            const int sl = n_img - n_first;
            const int V = B.meta[sl];
            const float* gt = B.gt[sl] + (size_t)p.h * m * 5;
            float maxiou = -INFINITY;                                                      // reduce_max over V = 0
            for (int j = 0; j < V; ++j) {
                GtBox t;
                t.x0 = gt[j * 5 + 0]; t.y0 = gt[j * 5 + 1]; t.x1 = gt[j * 5 + 2]; t.y1 = gt[j * 5 + 3];
                t.area = gt[j * 5 + 4];
                maxiou = fmaxf(maxiou, iou_ref(d.x0, d.y0, d.x1, d.y1, parea, t, g.tiou));   // :294
            }
            bool is_obj = false;                                                           // :328 scatter of owned anchors
            for (int i = 0; i < nl; ++i) is_obj |= (B.list_rec[i] == tid);
            const float bg = ((maxiou < thr) ? one : 0.0f) * (one - (is_obj ? one : 0.0f));   // :331-332
            const float pc = d.conf;
            float l = -logf(one - pc);                                                     // :335
            float fprime;
            if (g.focal) {
                const float pg = powf(pc, g.gamma);
                fprime = pg / (one - pc) - g.gamma * powf(pc, g.gamma - one) * logf(one - pc);
                l = l * pg;                                                                // :337
            } else {
                fprime = one / (one - pc);
            }
            const float contrib = l * bg;                                                  // :338 (0 * inf = NaN kept)
            const float sq = cur[0] * cur[0] + cur[1] * cur[1] + cur[2] * cur[2] + cur[3] * cur[3];   // :153-155
            if (p.h == 0) { acc_noobj[0] += contrib; acc_rect[0] += sq; }
            else if (p.h == 1) { acc_noobj[1] += contrib; acc_rect[1] += sq; }
            else { acc_noobj[2] += contrib; acc_rect[2] += sq; }
            if (with_grad) {
                const float pass = (d.sc >= g.lo && d.sc <= g.hi) ? one : 0.f;
                float gv[5];
                if (a.rect_active) {
                    const float rg = a.upstream * g.rw[p.h] * 2.0f;
                    gv[0] = rg * cur[0] / nf; gv[1] = rg * cur[1] / nf;
                    gv[2] = rg * cur[2] / nf; gv[3] = rg * cur[3] / nf;
                } else {
                    gv[0] = 0.f; gv[1] = 0.f; gv[2] = 0.f; gv[3] = 0.f;
                }
                gv[4] = a.kappa * bg * g.lw[p.h][2] * fprime * pass * d.sc * (one - d.sc);
                if (is_obj) {
                    // owner terms of every label assigned here, in label order (duplicates add, rule T3)
                    for (int i = 0; i < nl; ++i) {
                        if (B.list_rec[i] != tid) continue;
                        const OwnEntry* e = a.own + B.list_idx[i];
#pragma unroll
                        for (int c = 0; c < 5; ++c) gv[c] += e->g[c];
                    }
                }
                float* o = a.grad + (size_t)grec * L;
                if (a.rect_active || is_obj) {
#pragma unroll
                    for (int c = 0; c < 4; ++c) o[c] = gv[c];
                }
                o[4] = gv[4];
            }
        }

        // ---- records that own a label: class term and its gradient (one warp per record) ------------
        if (nl > 0 && C > 0) {
            for (int i = 0; i < nl; ++i) {
                const int rt = B.list_rec[i];
                if ((rt & (kWarps - 1)) != warp) continue;      // same record -> same warp, label order
                const OwnEntry* e = a.own + B.list_idx[i];
                const int cls = e->cls;
                if (cls < 0 || cls >= C) continue;               // zero one-hot row: no loss, no gradient
                const long long grec = rec0 + rt;
                const int h = locate(g, (int)(grec % g.rpi)).h;
                const float* z = a.predicts + (size_t)grec * L + 5;
                float mx = -INFINITY;
                for (int k = lane; k < C; k += 32) mx = fmaxf(mx, __ldg(z + k));
                mx = warp_max(mx);                                                         // decoder :189
                float sum = 0.f;
                for (int k = lane; k < C; k += 32) sum += expf(__ldg(z + k) - mx);
                sum = warp_sum(sum);
                const float inv = 1.0f / sum;
                const float qc = expf(__ldg(z + cls) - mx) * inv;                          // softmax :190
                const float qcc = clip_ref(qc, g.lo, g.hi);                                // :191
                if (lane == 0) acc_cls[h] += -logf(qcc);                                   // loss :363-364
                if (with_grad) {
                    const float pass = (qc >= g.lo && qc <= g.hi) ? 1.0f : 0.f;
                    const float coef = a.kappa * g.lw[h][4] * (-1.0f / qcc) * pass * qc;
                    float* go = a.grad + (size_t)grec * L + 5;
                    // on top of the zeros (ordered by the tile barrier); a second label on the same record
                    // is handled by the same lanes of the same warp, later in this loop
                    for (int k = lane; k < C; k += 32) {
                        const float qk = expf(__ldg(z + k) - mx) * inv;
                        go[k] = __ldcg(go + k) + coef * ((k == cls ? 1.0f : 0.f) - qk);
                    }
                }
                __syncwarp();
            }
        }
#pragma unroll
        for (int c = 0; c < 5; ++c) cur[c] = nxt[c];
    }

    // ---- reduction: thread -> warp -> CTA partial row -> last CTA --------------------------------
    float v[9] = {acc_noobj[0], acc_noobj[1], acc_noobj[2], acc_cls[0], acc_cls[1], acc_cls[2],
                  acc_rect[0], acc_rect[1], acc_rect[2]};
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const float w = warp_sum(v[k]);
        if (lane == 0) s.red[warp * 12 + k] = w;
    }
    __syncthreads();
    if (tid < 9) {
        float acc = 0.f;
        for (int w = 0; w < kWarps; ++w) acc += s.red[w * 12 + tid];
        a.partials[(size_t)blockIdx.x * kPartialFloats + tid] = acc;
    }
    __shared__ int is_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const int ticket = atomicAdd(&a.status[1], 1);
        is_last = (ticket == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    __shared__ float fin[18];
    if (tid < 18) {
        const int row = tid / 3, h = tid - row * 3;       // rows: xy, wh, noobj, obj, cls, rect
        double acc = 0.0;
        if (row == 0 || row == 1 || row == 3) {
            const int term = (row == 0) ? 0 : ((row == 1) ? 1 : 2);
            for (int i = 0; i < a.n; ++i) acc += (double)__ldcg(a.img_terms + (size_t)i * 9 + term * 3 + h);
        } else {
            const int col = (row == 2) ? h : ((row == 4) ? 3 + h : 6 + h);
            for (int b = 0; b < (int)gridDim.x; ++b) acc += (double)__ldcg(a.partials + (size_t)b * kPartialFloats + col);
            if (row == 2) acc *= (double)g.lw[h][2];
            else if (row == 4) acc *= (double)g.lw[h][4];
            else acc = a.rect_active ? acc * (double)g.rw[h] : 0.0;
        }
        const float out = (float)(acc / (double)a.n_global);
        a.detail[tid] = out;
        fin[tid] = out;
        if (!isfinite(out)) atomicOr((unsigned*)a.status, kStatusNonFinite);
    }
    __syncthreads();
    if (tid == 0) {
        float total = 0.f;                                // reduce_sum of the (6,3) detail (:137)
        for (int i = 0; i < 18; ++i) total += fin[i];
        a.detail[18] = total;
        a.detail[19] = 0.f;
        a.status[1] = 0;                                  // re-arm the ticket for the next launch
    }
}

cudaError_t launch_loss_dense(const Geo& g, DenseArgs a, int sm_count, cudaStream_t stream) {
    // a tile = one record per thread, never more than an image holds (a tile touches <= 2 images)
    long long tr = min((long long)kThreads, (long long)g.rpi);
    tr &= ~3LL;
    if (tr < 4) return cudaErrorInvalidValue;
    a.tr = (int)tr;
    a.num_tiles = (int)((a.total_rec + tr - 1) / tr);
    const size_t smem = dense_smem_bytes(a.m);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    cudaError_t err = cudaFuncSetAttribute(loss_dense_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 1;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, loss_dense_kernel, kThreads, smem);
    if (err != cudaSuccess) return err;
    if (per_sm < 1) per_sm = 1;
    const int grid = (int)min((long long)min(sm_count * per_sm, kMaxGrid), (long long)a.num_tiles);
    loss_dense_kernel<<<grid, kThreads, smem, stream>>>(g, a);
    return cudaGetLastError();
}

}  // namespace yv3
