// One sweep for both halves of an evaluation-while-training step: the score / threshold / rank pass of the
// post-process (postprocess.cu, score_tile_kernel) and the dense pass of the head loss (loss_dense.cu,
// loss_dense_kernel) on ONE read of the packed predicts.
//
// Why: run one after the other the two sweeps read the tensor twice -- the score sweep all of it (4*A*L bytes per
// image), the loss sweep a 128-byte DRAM line around every record's 20 bytes of box/conf logits -- and the loss sweep
// writes the gradient in 32-byte pieces around a bulk zero fill.  Here a warp lands a SPAN of S consecutive tiles (a
// tile = 32 records) in shared memory with one bulk (TMA) copy, and for each tile scores it (the same code:
// sweep_parts.cuh), runs the loss record phase on the five logits it already holds, then turns the tile IN PLACE into
// the gradient of those records; the whole span goes back with one bulk store: 4*A*L bytes in, 4*A*L bytes out, every
// DRAM line moved whole, once.
//
// Span length.  A tile is 128*L bytes: 10.9 KB at 80 classes, but 640 bytes at class_num 0 (the reference's shipped
// configuration, configs.py:42).  One tile per warp makes a class_num-0 sweep thousands of 640-byte copies, each with
// its own barrier, partial-sum reduction (45 shuffles) and partial row.  S grows as L shrinks (up to 4 tiles, S = 1
// from L = 64 up; see span_tiles for the measurements), but never so long that fewer than ~16 warps per SM are left.
// At L = 5 the sweep is NOT memory-bound: ~1000 instructions per record (three sigmoids, two exps, a log, the focal
// terms, up to 16 IoUs, ranking) against 40 bytes moved -- ncu shows the issue slots 85 % busy (profiles/).
//
// The arithmetic is the two kernels' own, expression for expression (yolov3_loss.py :267-294, :328-338, :140-164,
// :361-364; yolov3_post_process.py :44-76; the record phase and the owner phase are the SAME functions the dense
// kernel calls), so candidates, gradients and the integer decisions are bit-identical to the separate calls; the loss
// sums are folded per span instead of per CTA of eight warps (same fixed-order fold in double afterwards), which can
// move the last bit of a detail value.
#include "sweep_parts.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kMaxSpanTiles = 32;           // one object-bit word per tile, held by lane s

__global__ void __launch_bounds__(32, 24) fused_sweep_kernel(const __grid_constant__ Geo g, const __grid_constant__ DenseArgs a,
                                                         const ScoreTiles q, int S, float thresh, StageRow* __restrict__ stage,
                                                         int32_t* __restrict__ tcount, int32_t* __restrict__ post_status) {
    extern __shared__ __align__(128) unsigned char ssm[];
    __shared__ uint64_t bar_s;
    const int L = g.L, m = a.m;
    const int lane = threadIdx.x;
    const int span = blockIdx.x;
    const int tile0 = span * S;                                        // first tile (= chunk of 32 records) of the span
    const int ntile = min(S, q.num_tiles - tile0);
    float* buf = (float*)ssm;
    const long long srec0 = (long long)tile0 * 32;
    const int srec = (int)min((long long)ntile * 32, q.total_rec - srec0);
    const int sfl = srec * L;
    const unsigned bulk = (unsigned)(((size_t)sfl * 4) & ~(size_t)15);
    const float* src = a.predicts + (size_t)srec0 * L;
    if (bulk && lane == 0) {
        mbar_init(&bar_s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(&bar_s, bulk);
        bulk_g2s(buf, src, bulk, &bar_s);
    }
    // one object-bit word per tile (a tile starts on a multiple of 32 records): lane s holds tile s's
    const unsigned objw_mine = (lane < ntile) ? __ldg(a.objbits + tile0 + lane) : 0u;
    // the post-process status word starts clean with the sweep (the gather that may set bits in it comes after)
    if (span == 0 && lane < kStatusInts) post_status[lane] = 0;
    for (int i = (int)(bulk / 4) + lane; i < sfl; i += 32) buf[i] = __ldg(src + i);
    // where the span starts: one division per span, tiles step from there
    int n_first = (int)((unsigned)srec0 / (unsigned)g.rpi);
    int r_first = (int)((unsigned)srec0 - (unsigned)n_first * (unsigned)g.rpi);
    __syncwarp();                                      // barrier init visible to every lane
    if (bulk) mbar_wait(&bar_s, 0);
    __syncwarp();

    const float thr = g.iou_thresh;
    const float thr_lo = __fmul_rn(thr, 0.99999809265136719f);      // thr * (1 - 2^-19), see surely_below
    float an0 = 0.f, an1 = 0.f, an2 = 0.f;       // noobj sums per head
    float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f;       // sum of squared raw box logits per head
    float ac0 = 0.f, ac1 = 0.f, ac2 = 0.f;       // class sums per head
    const bool zero_fill = L > 5;                // at L = 5 the five values of every record cover the tile

    for (int s = 0; s < ntile; ++s) {
        float* tile = buf + (size_t)s * 32 * L;
        const int nrec = min(32, srec - s * 32);
        const bool ok = lane < nrec;
        const unsigned objw = __shfl_sync(0xffffffffu, objw_mine, s);
        const bool is_obj = ok && ((objw >> lane) & 1u);     // :328 scatter of owned anchors
        int n_img = n_first, r_img = r_first + lane, slot = 0;
        if (r_img >= g.rpi) { r_img -= g.rpi; ++n_img; slot = 1; }          // a tile touches at most two images
        float* r = tile + (size_t)lane * L;
        RecLoc p = {0, 0, 0, 0};
        float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f, sc = 0.f, conf = 0.f;
        if (ok) {
            p = locate(g, r_img);
            t0 = r[0]; t1 = r[1]; t2 = r[2]; t3 = r[3];
            sc = sigmoid_ref(r[4]);
            conf = clip_ref(sc, g.lo, g.hi);
        }
        // ---- post-process half: scores, threshold, ranks (reads the tile) ----------------------------------
        score_tile_records(g, q, tile, tile0 + s, lane, nrec, slot, p, conf, thresh, stage, tcount);

        // ---- loss half, record phase: lane = record ----------------------------------------------------------
        float gv0 = 0.f, gv1 = 0.f, gv2 = 0.f, gv3 = 0.f, gv4 = 0.f;
        if (ok) {
            const BoxDec d = decode_box(g, p, t0, t1, t2, t3, 0.f, true, sc);
            const float rc = (p.h == 0) ? a.rect_c[0] : ((p.h == 1) ? a.rect_c[1] : a.rect_c[2]);
            float noobj, sq;
            loss_record<false>(g, p, d, t0, t1, t2, t3, a.gtab + (size_t)n_img * gt_slot_floats(m), m, is_obj, true, a.kappa,
                               rc, a.rect_active, thr, thr_lo, gv0, gv1, gv2, gv3, gv4, noobj, sq);
            an0 += (p.h == 0) ? noobj : 0.f; an1 += (p.h == 1) ? noobj : 0.f; an2 += (p.h == 2) ? noobj : 0.f;
            ar0 += (p.h == 0) ? sq : 0.f; ar1 += (p.h == 1) ? sq : 0.f; ar2 += (p.h == 2) ? sq : 0.f;
        }

        // ---- the tile becomes the gradient of its records: zeros, then the five values of every record ------
        __syncwarp();                                      // every lane is done reading the tile
        if (zero_fill) {
            float4* t4v = reinterpret_cast<float4*>(tile);
            const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
            const int n16 = (nrec * L + 3) >> 2;           // the buffer is padded to 128 bytes
#pragma unroll 4
            for (int i = lane; i < n16; i += 32) t4v[i] = z4;
            __syncwarp();
        }
        if (ok) {
            r[0] = gv0; r[1] = gv1; r[2] = gv2; r[3] = gv3; r[4] = gv4;
        }
        // the next tile of the span
        r_first += 32;
        if (r_first >= g.rpi) { r_first -= g.rpi; ++n_first; }
    }
    fence_async_smem();
    __syncwarp();
    float* gdst = a.grad + (size_t)srec0 * L;
    if (lane == 0 && bulk) {
        bulk_s2g(gdst, buf, bulk);
        bulk_commit();
    }
    for (int i = (int)(bulk / 4) + lane; i < sfl; i += 32) gdst[i] = buf[i];

    // ---- records that own a label: the whole warp serves them one after the other (owner_record, shared with the
    // dense kernel), on top of the stored span; their five stored values are still in the shared-memory copy
    if (__any_sync(0xffffffffu, objw_mine != 0u) && !(a.debug & 4)) {
        if (lane == 0 && bulk) {
            bulk_wait_all<0>();                        // the span's store is complete, not only read
            asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        __syncwarp();
        for (int s = 0; s < ntile; ++s) {
            unsigned ob = __shfl_sync(0xffffffffu, objw_mine, s);
            const int nrec = min(32, srec - s * 32);
            if (nrec < 32) ob &= (1u << nrec) - 1u;
            while (ob) {
                const int srcl = __ffs(ob) - 1;
                ob &= ob - 1;
                const long long grec = srec0 + (long long)s * 32 + srcl;
                const int o_n = (int)((unsigned)grec / (unsigned)g.rpi);
                const int o_r = (int)((unsigned)grec - (unsigned)o_n * (unsigned)g.rpi);
                const int o_h = locate(g, o_r).h;
                const float gbase = (lane < 5) ? buf[((size_t)s * 32 + srcl) * L + lane] : 0.f;
                owner_record(g, a, o_n, o_r, o_h, (size_t)grec * L, gbase, true, lane, ac0, ac1, ac2);
            }
        }
    }

    // ---- loss sums: one partial row per span (loss_fold_kernel folds the rows in a fixed order) ------------------
    float v[9] = {an0, an1, an2, ac0, ac1, ac2, ar0, ar1, ar2};
    float mine = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const float w = warp_sum(v[k]);
        if (lane == k) mine = w;
    }
    if (lane < kPartialFloats) a.partials[((size_t)a.row_base + span) * kPartialFloats + lane] = (lane < 9) ? mine : 0.f;
    // the span must stay in place until the bulk store has read it
    if (lane == 0 && bulk) bulk_wait_read<0>();
}

// Tiles per span: ~16 KB of logits per warp, but at least ~16 warps per SM's worth of spans, and no more spans than
// partial rows.  0: the sweep does not apply.
static int span_tiles(const Geo& g, const ScoreTiles& q) {
    if (q.tr != 32 || !q.tma || q.num_tiles < 1) return 0;
    const long long tile_bytes = 128LL * g.L;
    // measured at class_num 0 (640-byte tiles) on B200: 88 / 72 / 68 / 67 / 68 us at S = 1 / 2 / 3 / 4 / 6 for batch 128, and
    // longer spans lose again (S = 16: 101 us): the tiles of a span are a serial chain inside one warp
    long long S = min(4LL, max(1LL, 16384 / tile_bytes));
    S = min(S, max(1LL, (long long)q.num_tiles / (148 * 16)));
    S = max(S, ((long long)q.num_tiles + kMaxGrid - 1) / kMaxGrid);
    if (const char* e = getenv("YV3_FUSED_SPAN")) {               // developer override
        const int v = atoi(e);
        if (v >= 1) S = v;
    }
    if (S > kMaxSpanTiles || S * tile_bytes > 200 * 1024) return 0;
    return (int)S;
}

int packed_tile_count(const Geo& g, const float* predicts, int n, float thresh) {
    return packed_score_tiles(g, predicts, n, thresh).num_tiles;
}

int fused_row_count(const Geo& g, const float* predicts, int n, float thresh) {
    const ScoreTiles q = packed_score_tiles(g, predicts, n, thresh);
    const int S = span_tiles(g, q);
    return S ? (q.num_tiles + S - 1) / S : 0;
}

// The fused sweep covers the packed, 16-byte aligned case with a gradient; everything else keeps the two sweeps.
bool fused_sweep_supported(const Geo& g, const DenseArgs& a, int n, float thresh) {
    if (!a.grad || !a.vec_ok) return false;
    const ScoreTiles q = packed_score_tiles(g, a.predicts, n, thresh);
    if (span_tiles(g, q) == 0) return false;
    if (a.debug & 32) return false;                    // developer switch: keep the two sweeps
    return true;
}

cudaError_t launch_fused_sweep(const Geo& g, DenseArgs a, int n, float thresh, StageRow* stage, int32_t* tcount,
                               int32_t* post_status, int* rows_out, cudaStream_t stream) {
    const ScoreTiles q = packed_score_tiles(g, a.predicts, n, thresh);
    const int S = span_tiles(g, q);
    if (!S) return cudaErrorInvalidValue;
    a.total_rec = q.total_rec;
    a.num_chunks = q.num_tiles;
    const int spans = (q.num_tiles + S - 1) / S;
    const size_t smem = align_up((size_t)S * 128 * g.L, 128);
    // the attribute sticks to the function on a device: raise it only when a launch needs more than any before
    static size_t smem_set[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || smem > smem_set[dev]) {
        cudaError_t err = cudaFuncSetAttribute(fused_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return err;
        if (dev >= 0 && dev < 64) smem_set[dev] = smem;
    }
    fused_sweep_kernel<<<spans, 32, smem, stream>>>(g, a, q, S, thresh, stage, tcount, post_status);
    if (rows_out) *rows_out = spans;
    return cudaGetLastError();
}

}  // namespace yv3
