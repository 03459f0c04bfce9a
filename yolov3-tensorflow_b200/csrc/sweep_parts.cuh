// Pieces shared by the score sweep (postprocess.cu), the dense loss sweep (loss_dense.cu) and the fused sweep that
// does both on one read of the tensor (fused_sweep.cu).
#pragma once
#include "tile_ops.cuh"

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
__device__ __forceinline__ void score_tile_records(const Geo& g, const ScoreTiles& q, const float* tile, int unit, int lane,
                                                   long long rec0, int nrec, float thresh, float4* __restrict__ stage,
                                                   int32_t* __restrict__ tcount) {
    const int L = g.L, C = g.C, tr = q.tr;
    const bool have = lane < nrec;
    const float* r = tile + (size_t)lane * L;
    float conf = 0.f;
    bool maybe = false;
    if (have) {
        conf = clip_ref(sigmoid_ref(r[4]), g.lo, g.hi);
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
    const int n_first = (int)(rec0 / g.rpi);
    int n_img = n_first, r_img = (int)(rec0 - (long long)n_first * g.rpi) + lane;
    if (r_img >= g.rpi) { r_img -= g.rpi; ++n_img; }
    RecLoc p = {0, 0, 0, 0};
    if (have) p = locate(g, r_img);
    const int grp = have ? (n_img - n_first) * 3 + p.h : 7;
    int my_rank = 0, my_base = 0, run = 0, cnt_lane = 0;
#pragma unroll
    for (int gi = 0; gi < 6; ++gi) {
        const unsigned b = __ballot_sync(0xffffffffu, pass && grp == gi);
        if (grp == gi) { my_rank = __popc(b & ((1u << lane) - 1u)); my_base = run; }
        if (lane == gi) cnt_lane = __popc(b);
        run += __popc(b);
    }
    if (pass) {
        const int rih = (p.y * g.W[p.h] + p.x) * g.B[p.h] + p.b;
        stage[(size_t)unit * tr + my_base + my_rank] = make_float4(mine.x, mine.y, mine.z, __int_as_float(rih));
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

}  // namespace yv3
