// Shared geometry + arithmetic of the head path (device side).
//
// Arithmetic follows the reference operator by operator, each operator rounded once to fp32:
//   decode   yolov3/yolov3_decoder.py:143-192      labels  yolov3/label_decoder.py:45-60
//   IoU      yolov3/yolov3_loss.py:267-299
// The explicit __f*_rn intrinsics are never contracted into FMAs (the file is also built with
// -fmad=false), so decisions taken on IoUs see the reference's roundings.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/yv3_head.h"

namespace yv3 {

constexpr int kHeads = 3;

// Everything a kernel needs to know about shapes and switches; built on the host from yv3_config
// (yv3_geometry() in yv3_api.cu) and passed by value as a __grid_constant__ kernel parameter.
struct Geo {
    int H[3], W[3], B[3];
    int fold[3];        // 16, 4, 1: head cells folded into one /32 row (yolov3_detector.py:80-85)
    int C, L;           // class_num, 5 + class_num
    int rows32;         // (H/32)*(W/32)
    int rpr;            // records per packed row  = 16*B8 + 4*B16 + B32
    int rpi;            // records per image       = rows32 * rpr
    int seg_base[3];    // first record of head h inside a packed row
    int head_rec[3];    // records of head h per image = H*W*B
    float aw[3][YV3_MAX_BOX], ah[3][YV3_MAX_BOX];   // fl(a) * fl(W_h), fl(a) * fl(H_h)  (decoder :35-40)
    float fW[3], fH[3], fHW[3];                      // float(W_h), float(H_h), float(H_h*W_h)
    float lo, hi;       // fp32(eps), fp32(1 - eps)
    float iou_thresh;
    float lw[3][5];     // xy, wh, noobj, obj, cls
    float rw[3];
    int focal;
    float alpha, gamma;
    int tiou;
    // ceil(2^32 / d) for the divisors of locate(): exact quotients for numerators < 2^22, d <= 2^10
    unsigned mg_rpr, mg_B[3], mg_W[3];      // 0 for d = 1
};

// The geometry of ONE head tensor (N, H, W, B*L) seen as a packed tensor that has only that head: the /32 row of
// yolov3_detector.py:80-85 holds fold*B records of it, starting at 0 (the Reshape alone, without the concatenate).
inline Geo head_geo(const Geo& g, int h) {
    Geo q = g;
    q.rpr = g.fold[h] * g.B[h];
    q.rpi = g.rows32 * q.rpr;
    const int far = 1 << 30;
    q.seg_base[0] = 0;
    q.seg_base[1] = (h >= 1) ? 0 : far;          // locate(): k >= seg_base[2] ? 2 : k >= seg_base[1] ? 1 : 0
    q.seg_base[2] = (h == 2) ? 0 : far;
    q.mg_rpr = q.rpr == 1 ? 0u : (unsigned)((1ULL << 32) / (unsigned long long)q.rpr + 1ULL);
    return q;
}

// x / d for x < 2^22 with mg = floor(2^32 / d) + 1, d <= 2^10 (x * (mg * d - 2^32) < 2^32)
__device__ __forceinline__ int fast_div(int x, unsigned mg) { return mg ? (int)__umulhi((unsigned)x, mg) : x; }

struct RecLoc {
    int h, y, x, b;
};

// record index inside an image (packed order) -> head / cell / anchor   
__device__ __forceinline__ RecLoc locate(const Geo& g, int r_img) {
    RecLoc o;
    int row = fast_div(r_img, g.mg_rpr);
    int k = r_img - row * g.rpr;
    o.h = (k >= g.seg_base[2]) ? 2 : ((k >= g.seg_base[1]) ? 1 : 0);
    int kk = k - g.seg_base[o.h];
    int B = g.B[o.h];
    int cell = fast_div(kk, g.mg_B[o.h]);
    o.b = kk - cell * B;
    int gidx = row * g.fold[o.h] + cell;
    int W = g.W[o.h];
    o.y = fast_div(gidx, g.mg_W[o.h]);
    o.x = gidx - o.y * W;
    return o;
}

// (head, row, col, anchor) -> record index inside an image
__device__ __forceinline__ int record_of(const Geo& g, int h, int y, int x, int b) {
    int gidx = y * g.W[h] + x;
    int row = gidx / g.fold[h];
    int cell = gidx - row * g.fold[h];
    return row * g.rpr + g.seg_base[h] + cell * g.B[h] + b;
}

// tf.sigmoid = 1 / (1 + exp(-t)): __frcp_rn is the correctly rounded reciprocal, i.e. the bits of __fdiv_rn(1, x),
// without the generic division's slow-path call
__device__ __forceinline__ float sigmoid_ref(float t) {
    return __frcp_rn(__fadd_rn(1.0f, expf(-t)));
}
// tf.pow(x, y) for the focal exponents (yolov3_loss.py:337, :346).  The reference's CPU kernel is std::pow, which
// returns the correctly rounded x * x for y = 2 (the shipped focal_gamma, configs.py:70) and x itself for y = 1;
// those two exponents (gamma and gamma - 1) take the one-instruction form here -- powf is ~100 instructions and
// no closer to the reference.  Any other exponent goes through powf.
__device__ __forceinline__ float pow_ref(float x, float y) {
    if (y == 2.0f) return __fmul_rn(x, x);
    if (y == 1.0f) return x;
    return powf(x, y);
}
__device__ __forceinline__ float clip_ref(float v, float lo, float hi) {
    return fmaxf(fminf(v, hi), lo);   // maximum(minimum(v, hi), lo): tf.clip_by_value
}

struct BoxDec {
    float sx, sy, sc;        // unclipped sigmoids of t0, t1, t4
    float ew, eh;            // exp(t2), exp(t3)
    float x, y, w, h, conf;  // decoded, grid units
    float x0, y0, x1, y1;    // corners
};

// sc: the unclipped sigmoid of t4 when the caller has it already (the score half of the fused sweep), else computed here
__device__ __forceinline__ BoxDec decode_box(const Geo& g, const RecLoc& p, float t0, float t1, float t2,
                                             float t3, float t4, bool have_sc = false, float sc = 0.f) {
    BoxDec d;
    d.sx = sigmoid_ref(t0);
    d.sy = sigmoid_ref(t1);
    d.sc = have_sc ? sc : sigmoid_ref(t4);
    d.x = __fadd_rn(clip_ref(d.sx, g.lo, g.hi), (float)p.x);
    d.y = __fadd_rn(clip_ref(d.sy, g.lo, g.hi), (float)p.y);
    d.ew = expf(t2);
    d.eh = expf(t3);
    d.w = __fmul_rn(d.ew, g.aw[p.h][p.b]);
    d.h = __fmul_rn(d.eh, g.ah[p.h][p.b]);
    d.conf = clip_ref(d.sc, g.lo, g.hi);
    float hw = __fmul_rn(d.w, 0.5f), hh = __fmul_rn(d.h, 0.5f);   // wh / 2 (exact)
    d.x0 = __fsub_rn(d.x, hw);
    d.y0 = __fsub_rn(d.y, hh);
    d.x1 = __fadd_rn(d.x, hw);
    d.y1 = __fadd_rn(d.y, hh);
    return d;
}

// One label in head-h grid units (label_decoder.py:53-59).
struct GtBox {
    float x0, y0, x1, y1, area;
};
struct GtDec {
    float tx, ty, tw, th;
    GtBox box;
};
__device__ __forceinline__ GtDec decode_label(const Geo& g, int h, float xn, float yn, float wn, float hn) {
    GtDec t;
    t.tx = __fmul_rn(xn, g.fW[h]);
    t.ty = __fmul_rn(yn, g.fH[h]);
    t.tw = __fmul_rn(wn, g.fW[h]);
    t.th = __fmul_rn(hn, g.fH[h]);
    float hw = __fmul_rn(t.tw, 0.5f), hh = __fmul_rn(t.th, 0.5f);
    t.box.x0 = __fsub_rn(t.tx, hw);
    t.box.y0 = __fsub_rn(t.ty, hh);
    t.box.x1 = __fadd_rn(t.tx, hw);
    t.box.y1 = __fadd_rn(t.ty, hh);
    t.box.area = __fmul_rn(t.tw, t.th);        // yolov3_loss.py:273
    return t;
}

// IoU exactly as yolov3_loss.py:276-293: inter / ((areaP + areaT) - inter) [* inter / areaT].
// A pair with an empty intersection is 0 / union = 0 and skips the divide.
__device__ __forceinline__ float iou_ref(float px0, float py0, float px1, float py1, float parea,
                                         const GtBox& t, int tiou) {
    float iw = fmaxf(__fsub_rn(fminf(px1, t.x1), fmaxf(px0, t.x0)), 0.0f);
    float ih = fmaxf(__fsub_rn(fminf(py1, t.y1), fmaxf(py0, t.y0)), 0.0f);
    float inter = __fmul_rn(iw, ih);
    if (inter == 0.0f) return 0.0f;
    float iou = __fdiv_rn(inter, __fsub_rn(__fadd_rn(parea, t.area), inter));
    if (tiou) iou = __fdiv_rn(__fmul_rn(iou, inter), t.area);
    return iou;
}

// ---- small PTX wrappers: mbarrier + 1-D bulk (TMA) copies ------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
// Bounded spin: a bulk copy that never lands traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
    uint32_t done = 0;
    for (unsigned spin = 0; spin < (1u << 26); ++spin) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}
// global -> shared, completes on an mbarrier (bytes: multiple of 16, both addresses 16-B aligned)
__device__ __forceinline__ void bulk_g2s(void* sdst, const void* gsrc, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global, tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* ssrc, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// make generic-proxy smem writes visible to the async proxy (TMA) before a bulk store reads them
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 2^x by the special-function unit alone (ex2.approx.ftz): relative error about 2^-22
__device__ __forceinline__ float exp2f_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---- workspace layout of the loss path (device memory handed in by the caller) ----------------
// [0]   int32 status[8]            : 0 error bits, 1 ticket counter of the final reduction
// then  uint32 objbits[ceil(n*rpi/32)] : bit r = record r (global index) is the anchor some owned label maps to
//                                    (the object mask of yolov3_loss.py:328); zeroed together with status
// then  OwnEntry own[n][3][m]      : assignment result, consumed by the dense pass
// then  float   img_terms[n][9]    : per image, per head weighted xy / wh / obj sums
// then  float   gtab[n][gt_slot_floats(m)] : per image [V, 3 pad], then per head the valid labels in head-h grid units,
//                                    label order kept: x0, y0, x1, y1, area, 3 pad  (label_decoder.py:53-59,
//                                    yolov3_loss.py:239-247, :273); see gt_slot_floats below
// then  int32   vcount[n]          : number of valid labels of an image
// then  float   partials[grid][12] : per CTA noobj[3], cls[3], rect[3] (see loss_dense.cu)
struct OwnEntry {
    int32_t rec;    // record index inside the image the label is assigned to in this head, or -1
    int32_t cls;    // int(label class); may lie outside [0, C)
    float g[5];     // owner contributions to d/dt0..t4, already scaled by upstream / n_global
    float pad;
};
static_assert(sizeof(OwnEntry) == 32, "OwnEntry must stay 32 bytes");

constexpr int kStatusInts = 8;
constexpr unsigned kStatusGtRange = 1u;
constexpr unsigned kStatusNonFinite = 2u;
constexpr unsigned kStatusCapacity = 4u;
constexpr int kGtFloats = 8;
// One image's label slot, the unit the dense pass copies to shared memory as it is: [V, pad x3], then per head label
// entries of 8 floats (x0, y0, x1, y1, area, pad x3), head stride m4*8 + 4 floats (m4 = m rounded up to 4) so that the
// three heads' entries fall into different banks.  Entries V .. round_up(V, 4) - 1 repeat label V - 1 (written by
// the assignment kernel): the background test walks the labels four at a time with constant offsets and no index
// clamping, and a repeated label does not change an "all below" decision.
__host__ __device__ inline int gt_head_stride(int m) { return ((m + 3) & ~3) * kGtFloats + 4; }
__host__ __device__ inline int gt_slot_floats(int m) { return 4 + 3 * gt_head_stride(m); }

struct Workspace {
    int32_t* status;
    uint32_t* objbits;
    size_t zero_bytes;      // status + objbits: one memset at the start of the assignment phase
    OwnEntry* own;
    float* img_terms;
    float* gtab;
    int32_t* vcount;
    float* partials;
    double* fold_slices;    // [64][9] first level of the two-level fold
    size_t bytes;
};
constexpr int kMaxGrid = 262144;        // rows of the partial-sum table: upper bound on CTAs of the dense pass / tiles of the fused sweep
constexpr int kPartialFloats = 12;      // noobj[3], cls[3], rect[3], pad[3]

__host__ __device__ inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
inline Workspace carve_workspace(void* base, int n, int m, long long rpi) {
    Workspace w;
    size_t off = 0;
    char* p = (char*)base;
    w.status = (int32_t*)(p + off);
    off = align_up(off + sizeof(int32_t) * kStatusInts, 32);
    w.objbits = (uint32_t*)(p + off);
    off = align_up(off + sizeof(uint32_t) * (size_t)(((long long)n * rpi + 31) / 32 + 8), 256);   // + slack: one region per head tensor in the un-packed form
    w.zero_bytes = off;
    w.own = (OwnEntry*)(p + off);
    off = align_up(off + sizeof(OwnEntry) * (size_t)n * 3 * (size_t)m, 256);
    w.img_terms = (float*)(p + off);
    off = align_up(off + sizeof(float) * 9 * (size_t)n, 256);
    w.gtab = (float*)(p + off);
    off = align_up(off + sizeof(float) * (size_t)gt_slot_floats(m) * (size_t)n, 256);
    w.vcount = (int32_t*)(p + off);
    off = align_up(off + sizeof(int32_t) * (size_t)n, 256);
    w.partials = (float*)(p + off);
    off = align_up(off + sizeof(float) * kPartialFloats * kMaxGrid, 256);
    w.fold_slices = (double*)(p + off);
    off = align_up(off + sizeof(double) * 9 * 64, 256);
    w.bytes = off;
    return w;
}

}  // namespace yv3
