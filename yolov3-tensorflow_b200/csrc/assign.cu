// Ground-truth -> cell / anchor / head assignment, and the owner-side loss terms.
//
// Follows yolov3/yolov3_loss.py: _get_valid_target :225-252 (valid <=> x * W8 >= 0), the response part
// of _calc_iou :269-271, :281, :296-302 (centre cell = floor(T_xy) reversed to (row, col); IoU of the
// label with the DECODED predictions of that cell's B anchors; max and first-index argmax), the
// cross-head `>=` ownership test :203-208, and for owned labels the obj / xy / wh terms of
// _single_head_loss :341-359 with the gradient TF autodiff derives for them.
//
// One CTA per image, one thread per (head, label slot).  The work is ~9 gathered records per label;
// the kernel is latency-bound and tiny next to the dense pass it feeds (loss_dense.cu).
#include "yv3_kernels.cuh"

namespace yv3 {


// (head, row, col, anchor) -> record index inside an image of the tensor that holds head h
__device__ __forceinline__ int rec_in_layout(const Geo& g, const HeadLayout& lay, int h, int y, int x, int b) {
    const int gidx = y * g.W[h] + x;
    const int row = gidx / g.fold[h];
    const int cell = gidx - row * g.fold[h];
    return row * lay.rpr[h] + lay.seg[h] + cell * g.B[h] + b;
}

__global__ void assign_kernel(const __grid_constant__ Geo g, const AssignArgs a) {
    extern __shared__ float sm[];
    const int m = a.m;
    float* rho_s = sm;                 // [3][m]
    float* term_s = sm + 3 * m;        // [3 terms][3][m]
    int* valid_s = (int*)(sm + 12 * m); // [m]
    const int n = blockIdx.x;
    const int tid = threadIdx.x;
    const bool active = tid < 3 * m;
    const int h = active ? tid / m : 0;
    const int j = active ? tid - h * m : 0;

    float xn = -1.f, yn = 0.f, wn = 0.f, hn = 0.f, cf = 0.f;
    if (active) {
        const float* t = a.targets + ((size_t)n * m + j) * 5;
        xn = __ldg(t + 0); yn = __ldg(t + 1); wn = __ldg(t + 2); hn = __ldg(t + 3); cf = __ldg(t + 4);
    }
    // padding test on the /8 head's x (yolov3_loss.py:239)
    bool valid = active && (__fmul_rn(xn, g.fW[0]) >= 0.0f);
    GtDec T = decode_label(g, h, xn, yn, wn, hn);
    int col = (int)floorf(T.tx), row = (int)floorf(T.ty);      // :269-270
    bool inside = valid && col >= 0 && col < g.W[h] && row >= 0 && row < g.H[h];
    if (valid && !inside) atomicOr((unsigned*)a.status, kStatusGtRange);

    float rho = 0.0f;
    int best = 0;
    BoxDec bd = {};
    if (inside) {
        RecLoc p; p.h = h; p.y = row; p.x = col; p.b = 0;
        const int rec0 = rec_in_layout(g, a.lay, h, row, col, 0);
        const float* base = a.lay.pred[h] + ((size_t)n * a.lay.rpi[h] + rec0) * g.L;
        rho = -INFINITY;
        for (int b = 0; b < g.B[h]; ++b) {
            const float* r = base + (size_t)b * g.L;
            float t0 = __ldg(r), t1 = __ldg(r + 1), t2 = __ldg(r + 2), t3 = __ldg(r + 3), t4 = __ldg(r + 4);
            p.b = b;
            BoxDec d = decode_box(g, p, t0, t1, t2, t3, t4);
            float parea = __fmul_rn(d.w, d.h);                          // :267
            float v = iou_ref(d.x0, d.y0, d.x1, d.y1, parea, T.box, g.tiou);
            if (b == 0 || v > rho) { rho = v; best = b; bd = d; }   // first max (:301)
        }
    }
    if (active) rho_s[h * m + j] = inside ? rho : 0.0f;
    if (active && h == 0) valid_s[j] = valid ? 1 : 0;
    __syncthreads();

    // the image's valid labels in this head's grid units, label order kept (:239-247): the list every anchor of
    // the image is compared with in the dense pass (:267-294)
    if (active) {
        int pos = 0, total = 0;
        for (int q = 0; q < m; ++q) {
            const int v = valid_s[q];
            pos += (q < j) ? v : 0;
            total += v;
        }
        if (valid) {
            float* slot = a.gtab + (size_t)n * gt_slot_floats(m);
            float4* o = reinterpret_cast<float4*>(slot + 4 + h * gt_head_stride(m)) + 2 * pos;
            const float4 bx = make_float4(T.box.x0, T.box.y0, T.box.x1, T.box.y1);
            const float4 ar = make_float4(T.box.area, 0.f, 0.f, 0.f);
            o[0] = bx;
            o[1] = ar;
            if (pos == total - 1) {                       // the last valid label pads the list to a multiple of four
                for (int e = total; e < ((total + 3) & ~3); ++e) {
                    o[2 * (e - pos)] = bx;
                    o[2 * (e - pos) + 1] = ar;
                }
            }
        }
        if (tid == 0) reinterpret_cast<int32_t*>(a.gtab + (size_t)n * gt_slot_floats(m))[0] = total;
        if (tid == 0) a.vcount[n] = total;
    }

    bool owned = false;
    if (inside) {
        float r1 = rho_s[((h + 1) % 3) * m + j], r2 = rho_s[((h + 2) % 3) * m + j];
        owned = (rho >= r1) && (rho >= r2);                              // :203-208
    }

    float t_xy = 0.f, t_wh = 0.f, t_obj = 0.f;
    OwnEntry e;
    e.rec = -1; e.cls = -1; e.pad = 0.f;
#pragma unroll
    for (int c = 0; c < 5; ++c) e.g[c] = 0.f;
    if (owned) {
        e.rec = rec_in_layout(g, a.lay, h, row, col, best);
        {
            const long long grec = (long long)n * a.lay.rpi[h] + e.rec;          // object mask bit (:328)
            atomicOr(a.lay.objbits[h] + (grec >> 5), 1u << (unsigned)(grec & 31));
        }
        e.cls = (int)cf;                                                 // tf.cast(., int32) truncates (:362)
        const float one = 1.0f;
        float sc = __fsub_rn(2.0f, __fdiv_rn(__fmul_rn(T.tw, T.th), g.fHW[h]));          // :350
        float kx = floorf(T.tx), ky = floorf(T.ty);                                       // :352
        float taux = __fsub_rn(T.tx, kx), tauy = __fsub_rn(T.ty, ky);                     // :353
        float ux = __fsub_rn(bd.x, kx), uy = __fsub_rn(bd.y, ky);                         // :354  (s + c) - c
        float lx = -(taux * logf(ux) + (one - taux) * logf(one - ux));                    // :355
        float ly = -(tauy * logf(uy) + (one - tauy) * logf(one - uy));
        t_xy = sc * lx + sc * ly;
        float dw = logf(T.tw) - logf(bd.w), dh = logf(T.th) - logf(bd.h);                 // :358
        t_wh = sc * (dw * dw) + sc * (dh * dh);
        float p = bd.conf;
        float lobj = -logf(p);                                                            // :344
        float gp;                                                                         // d lobj / d p
        if (g.focal) {
            float om = one - p;
            lobj = lobj * (pow_ref(om, g.gamma) * g.alpha);                               // :346
            gp = g.alpha * (-pow_ref(om, g.gamma) / p + g.gamma * pow_ref(om, g.gamma - one) * logf(p));
        } else {
            gp = -one / p;
        }
        t_obj = lobj;
        const float k = a.kappa;
        float pass_x = (bd.sx >= g.lo && bd.sx <= g.hi) ? one : 0.f;
        float pass_y = (bd.sy >= g.lo && bd.sy <= g.hi) ? one : 0.f;
        float pass_c = (bd.sc >= g.lo && bd.sc <= g.hi) ? one : 0.f;
        e.g[0] = k * g.lw[h][0] * sc * (-taux / ux + (one - taux) / (one - ux)) * pass_x * bd.sx * (one - bd.sx);
        e.g[1] = k * g.lw[h][0] * sc * (-tauy / uy + (one - tauy) / (one - uy)) * pass_y * bd.sy * (one - bd.sy);
        e.g[2] = k * g.lw[h][1] * sc * (-2.0f * dw / bd.w) * g.aw[h][best] * bd.ew;
        e.g[3] = k * g.lw[h][1] * sc * (-2.0f * dh / bd.h) * g.ah[h][best] * bd.eh;
        e.g[4] = k * g.lw[h][3] * gp * pass_c * bd.sc * (one - bd.sc);
    }
    if (active) {
        a.own[((size_t)n * 3 + h) * m + j] = e;
        term_s[(0 * 3 + h) * m + j] = t_xy;
        term_s[(1 * 3 + h) * m + j] = t_wh;
        term_s[(2 * 3 + h) * m + j] = t_obj;
        if (a.table) {
            int32_t* tb = a.table + (((size_t)n * m + j) * 3 + h) * 4;
            tb[0] = inside ? row : -1; tb[1] = inside ? col : -1; tb[2] = inside ? best : -1; tb[3] = owned ? 1 : 0;
        }
        if (a.iou_out) a.iou_out[((size_t)n * m + j) * 3 + h] = inside ? rho : 0.0f;
    }
    __syncthreads();
    // per image, per head sums in label order (fixed order => deterministic), then the head weight
    if (tid < 9) {
        int term = tid / 3, hh = tid - term * 3;
        const float* s = term_s + (term * 3 + hh) * m;
        float acc = 0.f;
        for (int q = 0; q < m; ++q) acc += s[q];
        const int widx = (term == 0) ? 0 : ((term == 1) ? 1 : 3);       // xy, wh, obj weights
        a.img_terms[(size_t)n * 9 + term * 3 + hh] = g.lw[hh][widx] * acc;
    }
}

cudaError_t launch_assign(const Geo& g, const AssignArgs& a_in, cudaStream_t stream) {
    AssignArgs a = a_in;
    if (!a.lay.pred[0]) {                                   // packed form: one tensor, one bit array
        for (int h = 0; h < 3; ++h) {
            a.lay.pred[h] = a.predicts;
            a.lay.rpi[h] = g.rpi; a.lay.rpr[h] = g.rpr; a.lay.seg[h] = g.seg_base[h];
            a.lay.objbits[h] = a.objbits;
        }
    }
    int threads = ((3 * a.m + 31) / 32) * 32;
    size_t smem = sizeof(float) * 13 * (size_t)a.m;
    assign_kernel<<<a.n, threads, smem, stream>>>(g, a);
    return cudaGetLastError();
}

}  // namespace yv3
