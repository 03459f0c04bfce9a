// Pieces shared by the score sweep (postprocess.cu), the dense loss sweep (loss_dense.cu) and the fused sweep that
// does both on one read of the tensor (fused_sweep.cu).
#pragma once
#include "tile_ops.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kSerialExactFrom = 6;     // exact scores: warp-cooperative per record up to this many per tile, per lane beyond

struct ScoreTiles {
    int tr;                 // records per tile = per WARP (multiple of 4, <= 32, <= records per image)
    int wpc;                // warps (tiles) per CTA
    int tile_floats;        // shared-memory floats reserved per tile (tr * L rounded up to 16 bytes)
    int num_tiles;
    long long total_rec;
    int tma;
    int early;              // threshold large enough for the approximate early-out (thresh > 2 * eps)
};

// the tiling of the packed sweep (postprocess.cu)
ScoreTiles packed_score_tiles(const Geo& g, const float* predicts, int n, float thresh);

// The score / threshold / rank part of the score sweep for the <= 32 records of one warp's tile, which is complete in
// shared memory when this is called (all 32 lanes call it): approximate early-out, exact scores of the queued
// records, ranking per (image slot, head), staged survivors and the six counts of the tile.
// `slot` (0 or 1: the tile's first or second image) and `p` locate the lane's record; `conf` is its clipped confidence
// clip(sigmoid(t4)) -- both are computed once by the caller, which needs them too.
__device__ __forceinline__ void score_tile_records(const Geo& g, const ScoreTiles& q, const float* tile, int unit, int lane,
                                                   int nrec, int slot, const RecLoc& p, float conf, float thresh,
                                                   StageRow* __restrict__ stage, int32_t* __restrict__ tcount) {
    const int L = g.L, C = g.C, tr = q.tr;
    const bool have = lane < nrec;
    const float* r = tile + (size_t)lane * L;
    bool maybe = false;
    if (have) {
        if (C == 0) {
            maybe = false;
        } else if (!q.early) {
            maybe = true;
        } else if (conf > thresh) {                                  // fl(prob * conf) <= conf for prob <= 1
            // ONE pass over the class logits: the sum is taken relative to the first logit c (S = sum 2^((z_k - c) log2e))
            // while the maximum M is tracked, then sum_k exp(z_k - M) = S * 2^(-(M - c) log2e).  The product of c
            // and log2e is the same rounded constant in every term, so it cancels; what is left is the rounding of
            // each exponent (|.| < 128, else a term is inf or 0 and the record goes to the exact path below).
            const float l2e = 1.4426950408889634f;
            const float z0 = r[5];
            const float cl = __fmul_rn(z0, l2e);
            float mx = z0;
            float s0 = 1.0f, s1 = 0.f, s2 = 0.f, s3 = 0.f;              // the first logit's own term is 2^0
            int k = 1;
            for (; k + 4 <= C; k += 4) {
                const float a0 = r[5 + k], a1 = r[6 + k], a2 = r[7 + k], a3 = r[8 + k];
                mx = fmaxf(fmaxf(mx, a0), a1);                         // FMNMX3
                mx = fmaxf(fmaxf(mx, a2), a3);
                s0 += exp2f_approx(__fmaf_rn(a0, l2e, -cl));
                s1 += exp2f_approx(__fmaf_rn(a1, l2e, -cl));
                s2 += exp2f_approx(__fmaf_rn(a2, l2e, -cl));
                s3 += exp2f_approx(__fmaf_rn(a3, l2e, -cl));
            }
            for (; k < C; ++k) {
                const float a0 = r[5 + k];
                mx = fmaxf(mx, a0);
                s0 += exp2f_approx(__fmaf_rn(a0, l2e, -cl));
            }
            const float S = (s0 + s1) + (s2 + s3);
            const float E = exp2f_approx(__fmaf_rn(-mx, l2e, cl));
            // true score <= conf / (S * E) * (1 + 1e-5); keep a 1e-3 band around the threshold for the exact path.
            // A sum that overflowed (or a NaN) proves nothing: both comparisons are written so that they fail then.
            maybe = !(conf < thresh * (S * E) * 0.999f) || !(S <= 3.0e38f);
        }
    }
    // ---- exact scores of the queued records: one record at a time, the classes spread over the warp's lanes ------
    float4 mine = make_float4(conf, 1.0f, __int_as_float(0), (have && conf > thresh) ? 1.f : 0.f);   // :53-55 (C = 0)
    if (C > 0) {
        mine = make_float4(0.f, 0.f, 0.f, 0.f);
        unsigned mb = __ballot_sync(0xffffffffu, maybe);
        if (__popc(mb) > kSerialExactFrom) {
            // many records of the tile need the exact score (a low threshold: nearly all of them): one record per lane,
            // the thread-serial class_stats -- bit-identical to the warp version by construction (tile_ops.cuh) -- costs
            // about six warp-cooperative records, whatever the count
            if (maybe) {
                const ClassStats cs = class_stats(r + 5, C);
                float best;
                int arg;
                class_best(r + 5, C, cs, g.lo, g.hi, best, arg);
                const float sc = __fmul_rn(best, conf);                                  // :59
                mine = make_float4(sc, best, __int_as_float(arg), (sc > thresh) ? 1.f : 0.f);   // :62 strict, fp32
            }
            mb = 0u;
        }
        while (mb) {
            const int src = __ffs(mb) - 1;
            mb &= mb - 1;
            const float* rr = tile + (size_t)src * L;
            const ClassStats cs = class_stats_warp(rr + 5, C, lane);
            float best = clip_ref(cs.inv, g.lo, g.hi);
            int arg = cs.arg;
            if (cs.near > 1) {                                                       // rounding tie: the serial rule
                if (lane == 0) class_best(rr + 5, C, cs, g.lo, g.hi, best, arg);
                best = __shfl_sync(0xffffffffu, best, 0);
                arg = __shfl_sync(0xffffffffu, arg, 0);
            }
            if (lane == src) {
                const float sc = __fmul_rn(best, conf);                              // :59
                mine = make_float4(sc, best, __int_as_float(arg), (sc > thresh) ? 1.f : 0.f);   // :62 strict, fp32
            }
        }
    }
    const bool pass = have && mine.w != 0.f;

    // ---- rank inside the tile per (image slot, head): ballots only -----------------------------------------
    const int grp = have ? slot * 3 + p.h : 7;
    int my_rank = 0, my_base = 0, run = 0, cnt_lane = 0;
#pragma unroll
    for (int gi = 0; gi < 6; ++gi) {
        const unsigned b = __ballot_sync(0xffffffffu, pass && grp == gi);
        if (grp == gi) { my_rank = __popc(b & ((1u << lane) - 1u)); my_base = run; }
        if (lane == gi) cnt_lane = __popc(b);
        run += __popc(b);
    }
    if (pass) {
        // the finished row (:65-75): the box is decoded here, from the logits the tile still holds
        const BoxDec d = decode_box(g, p, r[0], r[1], r[2], r[3], r[4]);
        StageRow o;
        o.box = make_float4(__fdiv_rn(d.x0, g.fW[p.h]), __fdiv_rn(d.y0, g.fH[p.h]), __fdiv_rn(d.x1, g.fW[p.h]),
                            __fdiv_rn(d.y1, g.fH[p.h]));                             // :65-68
        o.tail = make_float4(d.conf, mine.y, (float)__float_as_int(mine.z), mine.x);
        stage[(size_t)unit * tr + my_base + my_rank] = o;
    }
    if (lane < 6) tcount[(size_t)unit * 8 + lane] = cnt_lane;
}

// The background test of yolov3_loss.py:294,331 for one (anchor, label) pair: is IoU < iou_thresh?  The IoU itself
// is fl(inter / fl(fl(aP + aT) - inter)); the quotient is only needed when it lands close to the threshold.  With
// thr_lo = fl(thr * (1 - 2^-19)):   inter < fl(thr_lo * uni)   implies   inter / uni < thr * (1 - 2^-20)   (two
// roundings of at most 2^-24 each), hence fl(inter / uni) < thr.  `surely_below` is that one-multiply test; whatever
// it does not settle -- the band below the threshold, everything above it, NaNs, uni <= 0, thr <= 0, the TIoU variant
// -- takes the exact expression of iou_ref (`iou_below_exact`), so the decision is the reference's bit for bit.
__device__ __forceinline__ bool surely_below(float px0, float py0, float px1, float py1, float parea, const float4& t,
                                             float tarea, float thr_lo) {
    const float iw = fmaxf(__fsub_rn(fminf(px1, t.z), fmaxf(px0, t.x)), 0.0f);
    const float ih = fmaxf(__fsub_rn(fminf(py1, t.w), fmaxf(py0, t.y)), 0.0f);
    const float inter = __fmul_rn(iw, ih);
    const float uni = __fsub_rn(__fadd_rn(parea, tarea), inter);
    return inter < __fmul_rn(thr_lo, uni);
}
__device__ __forceinline__ bool iou_below_exact(float px0, float py0, float px1, float py1, float parea, const float4& t,
                                                float tarea, float thr, int tiou) {
    GtBox g;
    g.x0 = t.x; g.y0 = t.y; g.x1 = t.z; g.y1 = t.w; g.area = tarea;
    return iou_ref(px0, py0, px1, py1, parea, g, tiou) < thr;
}

// ---- the loss record phase: one thread = one anchor record (yolov3_loss.py :267-294, :328-338, :140-164) ----------
// Decoded box against the image's valid labels ("is every IoU below iou_thresh?", four labels at a time), the
// background mask, the noobj term (+ focal) and d/dt0..t4 of everything that reaches a record that owns no label.
// SHARED_GT: the label slot lies in shared memory (plain loads) instead of global memory (read-only path).
// Outputs: gv0..gv4 gradient values of t0..t4 (owner-side terms are added later, per label); noobj: this record's term of
// the noobj sum (:338; 0 * inf = NaN kept); sq: t0^2 + t1^2 + t2^2 + t3^2 (:153-155).
template <bool SHARED_GT>
__device__ __forceinline__ void loss_record(const Geo& g, const RecLoc& p, const BoxDec& d, float t0, float t1, float t2,
                                            float t3, const float* slot, int m, bool is_obj, bool with_grad,
                                            float kappa, float rc, int rect_active, float thr, float thr_lo,
                                            float& gv0, float& gv1, float& gv2, float& gv3, float& gv4, float& noobj, float& sq) {
    const float one = 1.0f;
    const float parea = __fmul_rn(d.w, d.h);                                       // :267
    const int V = SHARED_GT ? reinterpret_cast<const int*>(slot)[0] : __ldg(reinterpret_cast<const int*>(slot));
    const float4* gt = reinterpret_cast<const float4*>(slot + 4 + p.h * gt_head_stride(m));
    bool none_ge = true;                    // max over the labels < thr  <=>  every IoU < thr (V = 0: -inf < thr)
    for (int j = 0; j < V; j += 4) {
        // four labels in flight; the slots are padded to four (assign.cu), the last label repeated
        float4 bx[4];
        float ta[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (SHARED_GT) {
                bx[u] = gt[2 * (j + u)];
                ta[u] = reinterpret_cast<const float*>(gt + 2 * (j + u) + 1)[0];
            } else {
                bx[u] = __ldg(gt + 2 * (j + u));
                ta[u] = __ldg(reinterpret_cast<const float*>(gt + 2 * (j + u) + 1));
            }
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
        const float pg = pow_ref(pc, g.gamma);
        fprime = pg / (one - pc) - g.gamma * pow_ref(pc, g.gamma - one) * logf(one - pc);
        l = l * pg;                                                                // :337
    } else {
        fprime = one / (one - pc);
    }
    noobj = l * bg;                                                                // :338 (0 * inf = NaN kept)
    sq = t0 * t0 + t1 * t1 + t2 * t2 + t3 * t3;                                    // :153-155
    gv0 = gv1 = gv2 = gv3 = gv4 = 0.f;
    if (with_grad) {
        const float pass = (d.sc >= g.lo && d.sc <= g.hi) ? one : 0.f;
        if (rect_active) {
            // TF autodiff of rw_h * mean_N(sum t^2): rc = (upstream * rw_h) / N reaches every square, whose gradient
            // is dy * (2 t) (math_grad._SquareGrad); the quotient is per head and comes from the host
            gv0 = rc * (2.0f * t0); gv1 = rc * (2.0f * t1);
            gv2 = rc * (2.0f * t2); gv3 = rc * (2.0f * t3);
        }
        gv4 = kappa * bg * g.lw[p.h][2] * fprime * pass * d.sc * (one - d.sc);
        // TF's Exp gradient is dy * exp(t) (math_grad._ExpGrad) with a dense zero dy for anchors that own no label:
        // an overflowed exp(t2 / t3) turns it into 0 * inf = NaN, and so does the reference
        gv2 = __fadd_rn(gv2, __fmul_rn(0.f, d.ew));
        gv3 = __fadd_rn(gv3, __fmul_rn(0.f, d.eh));
    }
}

// ---- a record that owns labels (its object bit is set): served by the WHOLE WARP -----------------------------------
// The owner-side box/conf terms of assign.cu are added to the record's five values in label order (duplicates add,
// rule T3), then the class term (:361-364) and its softmax-CE gradient with the classes spread over the lanes, written
// on top of the gradient the sweep has already stored.  All 32 lanes call it; lanes 0..4 pass the stored value of
// channel `lane` in `gbase`.  The class loss lands in lane 0's ac0 / ac1 / ac2 (head o_h).
__device__ __forceinline__ void owner_record(const Geo& g, const DenseArgs& a, int o_n, int o_r, int o_h, size_t o_off,
                                             float gbase, bool with_grad, int lane, float& ac0, float& ac1, float& ac2) {
    const int C = g.C, m = a.m;
    const OwnEntry* e0 = a.own + ((size_t)o_n * 3 + o_h) * m;
    float mx = 0.f, inv = 0.f;
    bool have = false, first = true;
    float gacc = gbase;                                 // lanes 0..4: channel `lane` of the record
    bool touched = false;
    for (int jb = 0; jb < m; jb += 32) {
        const int je = jb + lane;
        const bool hit = je < m && e0[je].rec == o_r;
        unsigned hits = __ballot_sync(0xffffffffu, hit);
        while (hits) {                                 // label order
            const int j = jb + __ffs(hits) - 1;
            hits &= hits - 1;
            if (with_grad && lane < 5) { gacc += e0[j].g[lane]; touched = true; }
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
            if (with_grad) {
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
    }
    if (touched) a.grad[o_off + lane] = gacc;
}

}  // namespace yv3
