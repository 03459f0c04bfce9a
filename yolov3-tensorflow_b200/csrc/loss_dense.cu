// The fused, memory-bound pass of the head loss: decode + max-IoU + masks + noobj / rectified loss
// terms + the complete gradient tensor, one sweep over the packed predicts.
//
// Follows yolov3/yolov3_loss.py: the all-anchor part of _calc_iou :267-294, the masks of
// _single_head_loss :328-332, the noobj term :335-338, the class term :361-364, the rectified term
// :140-164, the batch mean :112 and the detail rows :115-138.  The backward is the closed form TF
// autodiff yields for those operators; no gradient flows through IoU or masks.
//
// Shape of the kernel (B200):
//  * the packed tensor is a flat array of records of L floats; a tile = TR consecutive records, whose
//    byte offset is a multiple of 16, so the gradient of a tile is ONE contiguous, 16-B aligned span;
//  * persistent CTAs (one per SM) walk tiles round-robin.  A thread owns one record of the tile (or
//    RPT of them when L is small): it reads the record's 5 box/conf logits (20 B out of L*4 -- class
//    logits of non-responsible anchors are never read, their gradient is identically 0), decodes,
//    takes the max IoU over the image's labels held in shared memory, and writes 5 gradient values
//    into a shared-memory image of the tile that is otherwise all zeros;
//  * labels of the <= 2 images a tile touches are staged in shared memory by a 1-D bulk (TMA) copy;
//  * the few records that own a label get their full gradient (box + softmax-CE) added by a warp;
//  * the finished tile leaves with one cp.async.bulk shared->global (TMA) store; two tile images
//    alternate so the store of tile i overlaps the arithmetic of tile i+1.  The zeros persist in
//    shared memory: only the class channels a fix-up dirtied are re-zeroed before reuse;
//  * loss sums: registers -> warp shuffles -> one partial row per CTA -> the last CTA to finish adds
//    the rows in index order (double) and writes the 18 detail values.  No float atomics anywhere.
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kSlots = 2;             // images a tile may touch (TR <= records per image)


struct DenseSmem {
    float* stage[2];                  // [tr * L]
    float* gt[kSlots];                // [3][m][5]  (x0, y0, x1, y1, area) compacted, label order
    float* lab_raw[kSlots];           // [m*5] raw label row (bulk-copy landing zone)
    int* gt_count;                    // [kSlots]
    int* list_rec;                    // [cap] tile-relative record of an owned entry
    int* list_idx;                    // [cap] index into own[]
    int* nlist;                       // [1]
    int* dirty[2];                    // [cap] per stage buffer: records whose class channels are dirty
    int* ndirty;                      // [2]
    unsigned char* objflag;           // [tr]
    float* red;                       // [kWarps][12]
    uint64_t* bar;                    // [kSlots] mbarriers of the label copies
};

__host__ __device__ inline size_t dense_smem_bytes(int tr, int L, int m, bool with_stage) {
    size_t b = 0;
    if (with_stage) b += 2 * align_up((size_t)tr * L * 4, 128);
    b += kSlots * align_up((size_t)3 * m * 5 * 4, 16);
    b += kSlots * align_up((size_t)m * 5 * 4, 16);
    b += 16;                                          // gt_count
    size_t cap = (size_t)kSlots * 3 * m;
    b += 2 * align_up(cap * 4, 16);                   // list_rec, list_idx
    b += 16;                                          // nlist
    b += 2 * align_up(cap * 4, 16);                   // dirty
    b += 16;                                          // ndirty
    b += align_up((size_t)tr, 16);                    // objflag
    b += kWarps * 12 * 4;
    b += 16 * kSlots;                                 // barriers
    return b + 128;
}

__device__ inline DenseSmem carve(unsigned char* base, int tr, int L, int m, bool with_stage) {
    DenseSmem s;
    size_t o = (128 - ((size_t)smem_u32(base) & 127)) & 127;
    auto take = [&](size_t bytes, size_t al) {
        o = (o + al - 1) / al * al;
        unsigned char* p = base + o;
        o += bytes;
        return p;
    };
    size_t stage_b = with_stage ? (size_t)tr * L * 4 : 0;
    s.stage[0] = (float*)take(stage_b, 128);
    s.stage[1] = (float*)take(stage_b, 128);
    for (int i = 0; i < kSlots; ++i) s.gt[i] = (float*)take((size_t)3 * m * 5 * 4, 16);
    for (int i = 0; i < kSlots; ++i) s.lab_raw[i] = (float*)take((size_t)m * 5 * 4, 16);
    s.gt_count = (int*)take(16, 16);
    size_t cap = (size_t)kSlots * 3 * m;
    s.list_rec = (int*)take(cap * 4, 16);
    s.list_idx = (int*)take(cap * 4, 16);
    s.nlist = (int*)take(16, 16);
    s.dirty[0] = (int*)take(cap * 4, 16);
    s.dirty[1] = (int*)take(cap * 4, 16);
    s.ndirty = (int*)take(16, 16);
    s.objflag = (unsigned char*)take((size_t)tr, 16);
    s.red = (float*)take(kWarps * 12 * 4, 16);
    s.bar = (uint64_t*)take(16 * kSlots, 16);
    return s;
}

template <int RPT>
__global__ void __launch_bounds__(kThreads, 1) loss_dense_kernel(const __grid_constant__ Geo g, const DenseArgs a) {
    extern __shared__ unsigned char smem_raw[];
    const bool with_stage = a.grad != nullptr;
    const DenseSmem s = carve(smem_raw, a.tr, g.L, a.m, with_stage);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int L = g.L, C = g.C, m = a.m, tr = a.tr;
    const int cap_entries = 3 * m;

    // ---- one-time init: zero tile images, flags, lists; arm the label barriers -----------------
    if (with_stage) {
        float4* z = (float4*)s.stage[0];
        const int n4 = (int)(((size_t)(s.stage[1] - s.stage[0]) * 4 + (size_t)tr * L * 4) / 16);
        for (int i = tid; i < n4; i += kThreads) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int i = tid; i < tr; i += kThreads) s.objflag[i] = 0;
    if (tid == 0) {
        s.nlist[0] = 0; s.ndirty[0] = 0; s.ndirty[1] = 0;
        s.gt_count[0] = 0; s.gt_count[1] = 0;
        for (int i = 0; i < kSlots; ++i) mbar_init(&s.bar[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    float acc_noobj[3] = {0.f, 0.f, 0.f}, acc_rect[3] = {0.f, 0.f, 0.f}, acc_cls[3] = {0.f, 0.f, 0.f};
    const float thr = g.iou_thresh;
    const float nf = (float)a.n_global;

    float cur[RPT][5];
    auto load_tile = [&](int tile, float (&dst)[RPT][5]) {
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            int ri = r * kThreads + tid;
            long long grec = (long long)tile * tr + ri;
            bool ok = tile < a.num_tiles && ri < tr && grec < a.total_rec;
            const float* p = a.predicts + (size_t)(ok ? grec : 0) * L;
#pragma unroll
            for (int c = 0; c < 5; ++c) dst[r][c] = ok ? __ldg(p + c) : 0.f;
        }
    };
    load_tile(blockIdx.x, cur);

    unsigned bar_phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        const long long rec0 = (long long)tile * tr;
        const int nrec = (int)min((long long)tr, a.total_rec - rec0);
        const int n_first = (int)(rec0 / g.rpi);
        const int n_last = (int)((rec0 + nrec - 1) / g.rpi);
        const int nslots = n_last - n_first + 1;          // 1 or 2

        // ---- tile setup ------------------------------------------------------------------------
        // (1) labels of the touched images -> shared memory (bulk copy when rows are 16-B aligned)
        if (a.tma_labels) {
            if (tid == 0) {
                fence_async_smem();       // earlier generic reads of lab_raw are ordered before the async writes
                for (int sl = 0; sl < nslots; ++sl) {
                    mbar_expect_tx(&s.bar[sl], (unsigned)(m * 20));
                    bulk_g2s(s.lab_raw[sl], a.targets + (size_t)(n_first + sl) * m * 5, (unsigned)(m * 20), &s.bar[sl]);
                }
            }
        } else {
            for (int i = tid; i < nslots * m * 5; i += kThreads) {
                int sl = i / (m * 5), k = i - sl * (m * 5);
                s.lab_raw[sl][k] = __ldg(a.targets + (size_t)(n_first + sl) * m * 5 + k);
            }
        }
        // (2) prefetch the next tile's logits while this one is processed
        float nxt[RPT][5];
        load_tile(tile + gridDim.x, nxt);
        // (3) the stage buffer must have left for global memory (store of two tiles ago)
        if (with_stage && tid == 0) bulk_wait_read<1>();
        if (a.tma_labels) {
            // each slot's barrier flips only in the tiles that use it: one phase bit per slot
            for (int sl = 0; sl < nslots; ++sl) mbar_wait(&s.bar[sl], (bar_phase >> sl) & 1u);
            bar_phase ^= (1u << nslots) - 1u;
        }
        __syncthreads();

        // (4) warp sl compacts image slot sl's valid labels (order kept) into per-head boxes;
        //     warp 2 collects the owned entries that fall into this tile (order kept)
        if (warp < nslots) {
            const int sl = warp;
            int count = 0;
            for (int base = 0; base < m; base += 32) {
                int j = base + lane;
                float xn = -1.f, yn = 0.f, wn = 0.f, hn = 0.f;
                if (j < m) {
                    const float* t = s.lab_raw[sl] + j * 5;
                    xn = t[0]; yn = t[1]; wn = t[2]; hn = t[3];
                }
                bool valid = (j < m) && (__fmul_rn(xn, g.fW[0]) >= 0.0f);
                unsigned mask = __ballot_sync(0xffffffffu, valid);
                int pos = count + __popc(mask & ((1u << lane) - 1u));
                if (valid) {
#pragma unroll
                    for (int h = 0; h < 3; ++h) {
                        GtDec T = decode_label(g, h, xn, yn, wn, hn);
                        float* o = s.gt[sl] + ((size_t)h * m + pos) * 5;
                        o[0] = T.box.x0; o[1] = T.box.y0; o[2] = T.box.x1; o[3] = T.box.y1; o[4] = T.box.area;
                    }
                }
                count += __popc(mask);
            }
            if (lane == 0) s.gt_count[sl] = count;
        } else if (warp == 2) {
            int count = 0;
            for (int sl = 0; sl < nslots; ++sl) {
                const int n_img = n_first + sl;
                const long long img_rec0 = (long long)n_img * g.rpi;
                for (int base = 0; base < cap_entries; base += 32) {
                    int e = base + lane;
                    int rec = -1;
                    if (e < cap_entries) rec = a.own[(size_t)n_img * cap_entries + e].rec;
                    long long grec = img_rec0 + rec;
                    bool in = rec >= 0 && grec >= rec0 && grec < rec0 + nrec;
                    unsigned mask = __ballot_sync(0xffffffffu, in);
                    int pos = count + __popc(mask & ((1u << lane) - 1u));
                    if (in) {
                        s.list_rec[pos] = (int)(grec - rec0);
                        s.list_idx[pos] = n_img * cap_entries + e;
                        s.objflag[(int)(grec - rec0)] = 1;
                    }
                    count += __popc(mask);
                }
            }
            if (lane == 0) s.nlist[0] = count;
        } else if (warp >= 3 && with_stage && C > 0) {
            // (5) re-zero the class channels the fix-ups of two tiles ago dirtied in this buffer
            const int nd = s.ndirty[buf];
            for (int d = warp - 3; d < nd; d += kWarps - 3) {
                float* rp = s.stage[buf] + (size_t)s.dirty[buf][d] * L + 5;
                for (int k = lane; k < C; k += 32) rp[k] = 0.f;
            }
        }
        __syncthreads();
        if (tid == 0) s.ndirty[buf] = 0;

        // ---- record phase ----------------------------------------------------------------------
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
            const int ri = r * kThreads + tid;
            if (ri < nrec) {
                const long long grec = rec0 + ri;
                const int n_img = (int)(grec / g.rpi);
                const int r_img = (int)(grec - (long long)n_img * g.rpi);
                const RecLoc p = locate(g, r_img);
                const BoxDec d = decode_box(g, p, cur[r][0], cur[r][1], cur[r][2], cur[r][3], cur[r][4]);
                const float parea = __fmul_rn(d.w, d.h);                                   // :267
                const int sl = n_img - n_first;
                const int V = s.gt_count[sl];
                const float* gt = s.gt[sl] + (size_t)p.h * m * 5;
                float maxiou = -INFINITY;                                                  // reduce_max over V = 0
                for (int j = 0; j < V; ++j) {
                    GtBox t;
                    t.x0 = gt[j * 5 + 0]; t.y0 = gt[j * 5 + 1]; t.x1 = gt[j * 5 + 2]; t.y1 = gt[j * 5 + 3];
                    t.area = gt[j * 5 + 4];
                    maxiou = fmaxf(maxiou, iou_ref(d.x0, d.y0, d.x1, d.y1, parea, t, g.tiou));   // :294
                }
                const float objf = s.objflag[ri] ? 1.0f : 0.0f;
                const float bg = ((maxiou < thr) ? 1.0f : 0.0f) * (1.0f - objf);           // :331-332
                const float pc = d.conf;
                const float one = 1.0f;
                float l = -logf(one - pc);                                                 // :335
                float fprime;
                if (g.focal) {
                    float pg = powf(pc, g.gamma);
                    fprime = pg / (one - pc) - g.gamma * powf(pc, g.gamma - one) * logf(one - pc);
                    l = l * pg;                                                            // :337
                } else {
                    fprime = one / (one - pc);
                }
                const float contrib = l * bg;                                              // :338 (0 * inf = NaN kept)
                const float sq = cur[r][0] * cur[r][0] + cur[r][1] * cur[r][1] + cur[r][2] * cur[r][2] +
                                 cur[r][3] * cur[r][3];                                    // :153-155
                if (p.h == 0) { acc_noobj[0] += contrib; acc_rect[0] += sq; }
                else if (p.h == 1) { acc_noobj[1] += contrib; acc_rect[1] += sq; }
                else { acc_noobj[2] += contrib; acc_rect[2] += sq; }
                if (with_stage) {
                    const float pass = (d.sc >= g.lo && d.sc <= g.hi) ? one : 0.f;
                    const float g4 = a.kappa * bg * g.lw[p.h][2] * fprime * pass * d.sc * (one - d.sc);
                    float* o = s.stage[buf] + (size_t)ri * L;
                    if (a.rect_active) {
                        const float rg = a.upstream * g.rw[p.h] * 2.0f;
                        o[0] = rg * cur[r][0] / nf; o[1] = rg * cur[r][1] / nf;
                        o[2] = rg * cur[r][2] / nf; o[3] = rg * cur[r][3] / nf;
                    } else {
                        o[0] = 0.f; o[1] = 0.f; o[2] = 0.f; o[3] = 0.f;
                    }
                    o[4] = g4;
                }
            }
        }
        __syncthreads();

        // ---- fix-up: records that own a label (one warp per record, entries in label order) ----
        const int nl = (a.debug & 4) ? 0 : s.nlist[0];
        for (int i = 0; i < nl; ++i) {
            const int rt = s.list_rec[i];
            if ((rt & (kWarps - 1)) != warp) continue;
            const OwnEntry* e = a.own + s.list_idx[i];
            const long long grec = rec0 + rt;
            const int r_img = (int)(grec % g.rpi);
            const int h = locate(g, r_img).h;
            float* o = with_stage ? s.stage[buf] + (size_t)rt * L : nullptr;
            if (with_stage && lane < 5) o[lane] += e->g[lane];
            if (C > 0) {
                const int cls = e->cls;
                const float* z = a.predicts + (size_t)grec * L + 5;
                float mx = -INFINITY;
                for (int k = lane; k < C; k += 32) mx = fmaxf(mx, __ldg(z + k));
                mx = warp_max(mx);                                                         // decoder :189
                float sum = 0.f;
                for (int k = lane; k < C; k += 32) sum += expf(__ldg(z + k) - mx);
                sum = warp_sum(sum);
                const float inv = 1.0f / sum;
                if (cls >= 0 && cls < C) {
                    const float qc = expf(__ldg(z + cls) - mx) * inv;                       // softmax :190
                    const float qcc = clip_ref(qc, g.lo, g.hi);                            // :191
                    if (lane == 0) acc_cls[h] += -logf(qcc);                               // loss :363-364
                    if (with_stage) {
                        const float pass = (qc >= g.lo && qc <= g.hi) ? 1.0f : 0.f;
                        const float coef = a.kappa * g.lw[h][4] * (-1.0f / qcc) * pass * qc;
                        for (int k = lane; k < C; k += 32) {
                            float qk = expf(__ldg(z + k) - mx) * inv;
                            o[5 + k] += coef * ((k == cls ? 1.0f : 0.f) - qk);
                        }
                    }
                }
                if (with_stage && lane == 0) {
                    int dpos = atomicAdd(&s.ndirty[buf], 1);
                    s.dirty[buf][dpos] = rt;
                }
            }
            __syncwarp();
        }
        __syncthreads();
        // reset the per-tile list state (warp 2 owns it)
        if (warp == 2) {
            for (int i = lane; i < nl; i += 32) s.objflag[s.list_rec[i]] = 0;
        }

        // ---- the tile leaves: one bulk shared->global store -------------------------------------
        if (with_stage) {
            fence_async_smem();
            __syncthreads();
            const size_t bytes = (size_t)nrec * L * 4;
            const unsigned bulk = (unsigned)(bytes & ~(size_t)15);
            float* gdst = a.grad + (size_t)rec0 * L;
            if (a.debug & 8) {
            } else if (a.debug & 2) {
                for (int i = tid; i < (int)(bytes / 4); i += kThreads) gdst[i] = s.stage[buf][i];
            } else {
            if (tid == 0) {
                if (bulk) bulk_s2g(gdst, s.stage[buf], bulk);
                bulk_commit();
            }
            const int tail = (int)((bytes - bulk) / 4);          // only the very last tile can have one
            if (tid < tail) gdst[bulk / 4 + tid] = s.stage[buf][bulk / 4 + tid];
            }
        }
#pragma unroll
        for (int r = 0; r < RPT; ++r)
#pragma unroll
            for (int c = 0; c < 5; ++c) cur[r][c] = nxt[r][c];
    }
    if (with_stage && tid == 0) bulk_wait_all<0>();

    // ---- reduction: thread -> warp -> CTA partial row -> last CTA --------------------------------
    float v[9] = {acc_noobj[0], acc_noobj[1], acc_noobj[2], acc_cls[0], acc_cls[1], acc_cls[2],
                  acc_rect[0], acc_rect[1], acc_rect[2]};
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        float w = warp_sum(v[k]);
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
        int ticket = atomicAdd(&a.status[1], 1);
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
    const bool with_stage = a.grad != nullptr;
    // records per tile: as many as fit two tile images next to the label / list areas in the 227 KB
    // of shared memory a CTA may own, at most 8 per thread
    const size_t fixed = dense_smem_bytes(kThreads * 8, g.L, a.m, false);
    const size_t limit = 232448 - 1024;
    if (fixed + 64 * (size_t)g.L > limit) return cudaErrorInvalidValue;
    const size_t budget = with_stage ? ((limit - fixed) / 2 - 128) : ((size_t)1 << 30);
    long long tr = (long long)(budget / ((size_t)g.L * 4));
    tr = min(tr, (long long)kThreads * 8);
    tr = min(tr, (long long)g.rpi);
    tr &= ~3LL;
    if (tr < 4) return cudaErrorInvalidValue;
    if (tr > kThreads) tr = tr / kThreads * kThreads;      // whole records-per-thread
    a.tr = (int)tr;
    a.num_tiles = (int)((a.total_rec + tr - 1) / tr);
    const int rpt = (int)((tr + kThreads - 1) / kThreads);
    const int grid = (int)min((long long)min(sm_count, kMaxGrid), (long long)a.num_tiles);
    const size_t smem = dense_smem_bytes(a.tr, g.L, a.m, with_stage);
    cudaError_t err = cudaSuccess;
#define YV3_LAUNCH(R)                                                                                    \
    do {                                                                                                 \
        err = cudaFuncSetAttribute(loss_dense_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                   (int)smem);                                                           \
        if (err != cudaSuccess) return err;                                                              \
        loss_dense_kernel<R><<<grid, kThreads, smem, stream>>>(g, a);                                    \
    } while (0)
    switch (rpt) {
        case 1: YV3_LAUNCH(1); break;
        case 2: YV3_LAUNCH(2); break;
        case 3: YV3_LAUNCH(3); break;
        case 4: YV3_LAUNCH(4); break;
        case 5: case 6: case 7: case 8: YV3_LAUNCH(8); break;
        default: return cudaErrorInvalidValue;
    }
#undef YV3_LAUNCH
    return cudaGetLastError();
}

}  // namespace yv3
