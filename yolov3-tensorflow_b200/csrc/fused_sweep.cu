// One sweep for both halves of an evaluation-while-training step: the score / threshold / rank pass of the
// post-process (postprocess.cu, score_tile_kernel) and the dense pass of the head loss (loss_dense.cu,
// loss_dense_kernel) on ONE read of the packed predicts.
//
// Why: run one after the other the two sweeps read the tensor twice -- the score sweep all of it (4*A*L bytes per
// image), the loss sweep a 128-byte DRAM line around every record's 20 bytes of box/conf logits -- and the loss sweep
// writes the gradient in 32-byte pieces around a bulk zero fill.  Here a warp lands its tile of 32 records in shared
// memory with one bulk (TMA) copy, scores it (the same code: sweep_parts.cuh), runs the loss record phase on the five
// logits it already holds, then turns the tile IN PLACE into the gradient of those records (zeros, the five values per
// record) and sends it back with one bulk store: 4*A*L bytes in, 4*A*L bytes out, every DRAM line moved whole, once.
//
// The arithmetic is the two kernels' own, expression for expression (yolov3_loss.py :267-294, :328-338, :140-164,
// :361-364; yolov3_post_process.py :44-76), so candidates, gradients and the integer decisions are bit-identical to
// the separate calls; the loss sums are folded per warp instead of per CTA of eight warps (same fixed-order fold in
// double afterwards), which can move the last bit of a detail value.
#include "sweep_parts.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

__global__ void __launch_bounds__(32) fused_sweep_kernel(const __grid_constant__ Geo g, const __grid_constant__ DenseArgs a,
                                                         const ScoreTiles q, float thresh, float4* __restrict__ stage,
                                                         int32_t* __restrict__ tcount, int32_t* __restrict__ post_status) {
    extern __shared__ __align__(128) unsigned char ssm[];
    __shared__ uint64_t bar_s;
    const int L = g.L, C = g.C, m = a.m;
    const int lane = threadIdx.x;
    const int unit = blockIdx.x;                       // tile = chunk of 32 records (q.tr == 32)
    float* tile = (float*)ssm;
    const long long rec0 = (long long)unit * 32;
    const int nrec = (int)min(32LL, q.total_rec - rec0);
    const int nfl = nrec * L;
    const unsigned bulk = (unsigned)(((size_t)nfl * 4) & ~(size_t)15);
    const float* src = a.predicts + (size_t)rec0 * L;
    if (bulk) {
        if (lane == 0) {
            mbar_init(&bar_s, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            mbar_expect_tx(&bar_s, bulk);
            bulk_g2s(tile, src, bulk, &bar_s);
        }
    }
    const unsigned objw = __ldg(a.objbits + unit);     // rec0 is a multiple of 32: one word per tile
    // the post-process status word starts clean with the sweep (the gather that may set bits in it comes after)
    if (unit == 0 && lane < kStatusInts) post_status[lane] = 0;
    for (int i = (int)(bulk / 4) + lane; i < nfl; i += 32) tile[i] = __ldg(src + i);
    __syncwarp();                                      // barrier init visible to every lane
    if (bulk) mbar_wait(&bar_s, 0);
    __syncwarp();

    // ---- post-process half: scores, threshold, ranks (reads the tile) --------------------------------------
    score_tile_records(g, q, tile, unit, lane, rec0, nrec, thresh, stage, tcount);

    // ---- loss half, record phase: lane = record (loss_dense_kernel's, logits from the tile) -----------------
    const float thr = g.iou_thresh;
    const float thr_lo = __fmul_rn(thr, 0.99999809265136719f);      // thr * (1 - 2^-19), see surely_below
    const float nf = (float)a.n_global;
    const float one = 1.0f;
    float an0 = 0.f, an1 = 0.f, an2 = 0.f;       // noobj sums per head
    float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f;       // sum of squared raw box logits per head
    float ac0 = 0.f, ac1 = 0.f, ac2 = 0.f;       // class sums per head
    const bool ok = lane < nrec;
    const bool is_obj = ok && ((objw >> lane) & 1u);     // :328 scatter of owned anchors
    float gv[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    const int n_first = (int)((unsigned)rec0 / (unsigned)g.rpi);
    const int r_first = (int)((unsigned)rec0 - (unsigned)n_first * (unsigned)g.rpi);
    int n_img = n_first, r_img = r_first + lane, head = 0;
    float* r = tile + (size_t)lane * L;
    if (ok) {
        const float t0 = r[0], t1 = r[1], t2 = r[2], t3 = r[3], t4 = r[4];
        while (r_img >= g.rpi) { r_img -= g.rpi; ++n_img; }
        const RecLoc p = locate(g, r_img);
        head = p.h;
        const BoxDec d = decode_box(g, p, t0, t1, t2, t3, t4);
        const float parea = __fmul_rn(d.w, d.h);                                       // :267
        const float* slot = a.gtab + (size_t)n_img * gt_slot_floats(m);
        const int V = __ldg(reinterpret_cast<const int*>(slot));
        const float4* gt = reinterpret_cast<const float4*>(slot + 4 + p.h * gt_head_stride(m));
        bool none_ge = true;                    // max over the labels < thr  <=>  every IoU < thr (V = 0: -inf < thr)
        for (int j = 0; j < V; j += 4) {
            // four labels in flight; the slots are padded to four (assign.cu), the last label repeated
            float4 bx[4];
            float ta[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                bx[u] = __ldg(gt + 2 * (j + u));
                ta[u] = __ldg(reinterpret_cast<const float*>(gt + 2 * (j + u) + 1));
            }
            bool sure = !g.tiou;
#pragma unroll
            for (int u = 0; u < 4; ++u) sure = sure & surely_below(d.x0, d.y0, d.x1, d.y1, parea, bx[u], ta[u], thr_lo);
            if (!sure) {                                                       // rare: some IoU near or above thr
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    none_ge = none_ge && iou_below_exact(d.x0, d.y0, d.x1, d.y1, parea, bx[u], ta[u], thr, g.tiou);   // :294
            }
        }
        const float bg = (none_ge ? one : 0.0f) * (one - (is_obj ? one : 0.0f));       // :331-332
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
        const float sq = t0 * t0 + t1 * t1 + t2 * t2 + t3 * t3;                        // :153-155
        an0 += (p.h == 0) ? contrib : 0.f; an1 += (p.h == 1) ? contrib : 0.f; an2 += (p.h == 2) ? contrib : 0.f;
        ar0 += (p.h == 0) ? sq : 0.f; ar1 += (p.h == 1) ? sq : 0.f; ar2 += (p.h == 2) ? sq : 0.f;
        const float pass = (d.sc >= g.lo && d.sc <= g.hi) ? one : 0.f;
        if (a.rect_active) {
            const float rg = a.upstream * g.rw[p.h] * 2.0f;
            gv[0] = rg * t0 / nf; gv[1] = rg * t1 / nf;
            gv[2] = rg * t2 / nf; gv[3] = rg * t3 / nf;
        }
        gv[4] = a.kappa * bg * g.lw[p.h][2] * fprime * pass * d.sc * (one - d.sc);
        // TF's Exp gradient is dy * exp(t) (math_grad._ExpGrad) with a dense zero dy for anchors that own no label:
        // an overflowed exp(t2 / t3) turns it into 0 * inf = NaN, and so does the reference
        gv[2] = __fadd_rn(gv[2], __fmul_rn(0.f, d.ew));
        gv[3] = __fadd_rn(gv[3], __fmul_rn(0.f, d.eh));
    }

    // ---- the tile becomes the gradient of its records: zeros, then the five values of every record ----------
    __syncwarp();                                      // every lane is done reading the tile
    {
        float4* t4v = reinterpret_cast<float4*>(tile);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const int n16 = (nfl + 3) >> 2;                // the buffer is padded to 128 bytes
#pragma unroll 4
        for (int i = lane; i < n16; i += 32) t4v[i] = z4;
    }
    __syncwarp();
    if (ok) {
        r[0] = gv[0]; r[1] = gv[1]; r[2] = gv[2]; r[3] = gv[3]; r[4] = gv[4];
    }
    fence_async_smem();
    __syncwarp();
    float* gdst = a.grad + (size_t)rec0 * L;
    if (lane == 0 && bulk) {
        bulk_s2g(gdst, tile, bulk);
        bulk_commit();
    }
    for (int i = (int)(bulk / 4) + lane; i < nfl; i += 32) gdst[i] = tile[i];

    // ---- records that own a label: the whole warp serves them one after the other (loss_dense_kernel's owner
    // phase: owner-side box/conf terms of assign.cu added in label order, duplicates add, rule T3; then the class
    // term :361-364 and its softmax-CE gradient with the classes spread over the lanes), on top of the stored tile
    unsigned ob = __ballot_sync(0xffffffffu, is_obj);
    if (ob && !(a.debug & 4)) {
        if (lane == 0) {
            bulk_wait_all<0>();                        // the tile's store is complete, not only read
            asm volatile("fence.proxy.async.global;" ::: "memory");
        }
        __syncwarp();
        while (ob) {
            const int srcl = __ffs(ob) - 1;
            ob &= ob - 1;
            const int o_n = __shfl_sync(0xffffffffu, n_img, srcl);
            const int o_r = __shfl_sync(0xffffffffu, r_img, srcl);
            const int o_h = __shfl_sync(0xffffffffu, head, srcl);
            const size_t o_off = (size_t)(rec0 + srcl) * L;
            const OwnEntry* e0 = a.own + ((size_t)o_n * 3 + o_h) * m;
            float mx = 0.f, inv = 0.f;
            bool have = false, first = true;
            float gacc = 0.f;                                   // lanes 0..4: channel `lane` of the record
#pragma unroll
            for (int c = 0; c < 5; ++c) {
                const float v = __shfl_sync(0xffffffffu, gv[c], srcl);
                if (lane == c) gacc = v;
            }
            bool touched = false;
            for (int jb = 0; jb < m; jb += 32) {
                const int je = jb + lane;
                const bool hit = je < m && e0[je].rec == o_r;
                unsigned hits = __ballot_sync(0xffffffffu, hit);
                while (hits) {                                 // label order
                    const int j = jb + __ffs(hits) - 1;
                    hits &= hits - 1;
                    if (lane < 5) { gacc += e0[j].g[lane]; touched = true; }
                    const int cls = e0[j].cls;
                    if (C == 0 || cls < 0 || cls >= C) continue;     // zero one-hot row: no loss, no gradient
                    const float* z = a.predicts + o_off + 5;
                    if (!have) {                               // decoder :189-190
                        float vmax = -INFINITY;
                        for (int k = lane; k < C; k += 32) vmax = fmaxf(vmax, __ldg(z + k));
                        mx = warp_max(vmax);
                        float sum = 0.f;
                        for (int k = lane; k < C; k += 32) sum += expf(__ldg(z + k) - mx);
                        inv = 1.0f / warp_sum(sum);
                        have = true;
                    }
                    const float qc = expf(__ldg(z + cls) - mx) * inv;
                    const float qcc = clip_ref(qc, g.lo, g.hi);                                // :191
                    if (lane == 0) {
                        const float lv = -logf(qcc);                                           // loss :363-364
                        ac0 += (o_h == 0) ? lv : 0.f; ac1 += (o_h == 1) ? lv : 0.f; ac2 += (o_h == 2) ? lv : 0.f;
                    }
                    const float pass = (qc >= g.lo && qc <= g.hi) ? 1.0f : 0.f;
                    const float coef = a.kappa * g.lw[o_h][4] * (-1.0f / qcc) * pass * qc;
                    float* go = a.grad + o_off + 5;
                    // a second label on the same record: the same lane adds to what it wrote for the first
                    for (int k = lane; k < C; k += 32) {
                        const float qk = expf(__ldg(z + k) - mx) * inv;
                        const float v = coef * ((k == cls ? 1.0f : 0.f) - qk);
                        go[k] = first ? v : (__ldcg(go + k) + v);
                    }
                    first = false;
                }
            }
            if (touched) a.grad[o_off + lane] = gacc;
        }
    }

    // ---- loss sums: one partial row per warp (loss_fold_kernel folds the rows in a fixed order) --------------
    float v[9] = {an0, an1, an2, ac0, ac1, ac2, ar0, ar1, ar2};
    float mine = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const float w = warp_sum(v[k]);
        if (lane == k) mine = w;
    }
    if (lane < kPartialFloats) a.partials[((size_t)a.row_base + unit) * kPartialFloats + lane] = (lane < 9) ? mine : 0.f;
    // the tile must stay in place until the bulk store has read it
    if (lane == 0 && bulk) bulk_wait_read<0>();
}

int packed_tile_count(const Geo& g, const float* predicts, int n, float thresh) {
    return packed_score_tiles(g, predicts, n, thresh).num_tiles;
}

// The fused sweep covers the packed, 16-byte aligned case with a gradient; everything else keeps the two sweeps.
bool fused_sweep_supported(const Geo& g, const DenseArgs& a, int n, float thresh) {
    if (!a.grad || !a.vec_ok) return false;
    const ScoreTiles q = packed_score_tiles(g, a.predicts, n, thresh);
    if (q.tr != 32 || !q.tma) return false;
    if (q.num_tiles > kMaxGrid) return false;          // one partial row per tile
    if ((size_t)q.tile_floats * 4 > 200 * 1024) return false;
    if (a.debug & 32) return false;                    // developer switch: keep the two sweeps
    return true;
}

cudaError_t launch_fused_sweep(const Geo& g, DenseArgs a, int n, float thresh, float4* stage, int32_t* tcount,
                               int32_t* post_status, int* rows_out, cudaStream_t stream) {
    const ScoreTiles q = packed_score_tiles(g, a.predicts, n, thresh);
    a.total_rec = q.total_rec;
    a.num_chunks = q.num_tiles;
    const size_t smem = (size_t)q.tile_floats * 4;
    cudaError_t err = cudaFuncSetAttribute(fused_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    fused_sweep_kernel<<<q.num_tiles, 32, smem, stream>>>(g, a, q, thresh, stage, tcount, post_status);
    if (rows_out) *rows_out = q.num_tiles;
    return cudaGetLastError();
}

}  // namespace yv3
