/* TEST INFRASTRUCTURE ONLY (oracle). Not linked into the product.
 *
 * Plain-C twin of oracle/ref_postprocess.py::greedy_nms, for candidate counts where the
 * reference's O(K^2) pure-Python loop (yolov3/yolov3_post_process.py:108-129) takes minutes.
 * Same rules: stable sort by score descending (:115), suppress a later box iff class ids compare
 * equal as doubles (:119) and IoU > thresh strictly (:124); IoU in float64 with the reference's
 * operation order and early-out (:139-145).  Build with -ffp-contract=off (see oracle/Makefile).
 * Parity: pinned through tests/golden/postprocess_*.npz (outputs of the reference itself).
 */
#include <stdint.h>
#include <stdlib.h>

static const double *g_score;

static int cmp_desc_stable(const void *a, const void *b) {
    int64_t ia = *(const int64_t *)a, ib = *(const int64_t *)b;
    double sa = g_score[ia * 8 + 7], sb = g_score[ib * 8 + 7];
    if (sa > sb) return -1;
    if (sa < sb) return 1;
    return (ia > ib) - (ia < ib);
}

static double iou_f64(const double *box, const double *truth) {
    double l = box[0] > truth[0] ? box[0] : truth[0];
    double r = box[2] < truth[2] ? box[2] : truth[2];
    double w = r - l;
    double t = box[1] > truth[1] ? box[1] : truth[1];
    double b = box[3] < truth[3] ? box[3] : truth[3];
    double h = b - t;
    if (w <= 0 || h <= 0) return 0.0;
    double inter = w * h;
    double uni = (box[2] - box[0]) * (box[3] - box[1]) + (truth[2] - truth[0]) * (truth[3] - truth[1]) - inter;
    return inter / uni;
}

/* cands: K rows of 8 doubles [x0,y0,x1,y1,conf,prob,cls,score] in pre-sort order.
 * keep_out: receives indices (into cands) of survivors in sorted order; returns their count. */
int64_t yv3o_greedy_nms(const double *cands, int64_t K, double thresh, int64_t *keep_out) {
    if (K <= 0) return 0;
    int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (size_t)K);
    unsigned char *dead = (unsigned char *)calloc((size_t)K, 1);
    for (int64_t i = 0; i < K; ++i) order[i] = i;
    g_score = cands;
    qsort(order, (size_t)K, sizeof(int64_t), cmp_desc_stable);
    int64_t nk = 0;
    for (int64_t oi = 0; oi < K; ++oi) {
        int64_t i = order[oi];
        if (dead[i]) continue;
        keep_out[nk++] = i;
        const double *bi = cands + i * 8;
        for (int64_t oj = oi + 1; oj < K; ++oj) {
            int64_t j = order[oj];
            if (dead[j]) continue;
            const double *bj = cands + j * 8;
            if (bj[6] != bi[6]) continue;
            if (iou_f64(bi, bj) > thresh) dead[j] = 1;
        }
    }
    free(order);
    free(dead);
    return nk;
}
