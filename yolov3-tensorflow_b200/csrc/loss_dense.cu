// The fused, memory-bound pass of the head loss: decode + max-IoU + masks + noobj / rectified / class
// loss terms + the complete gradient tensor, one sweep over the packed predicts.
//
// Follows yolov3/yolov3_loss.py: the all-anchor part of _calc_iou :267-294, the masks of
// _single_head_loss :328-332, the noobj term :335-338, the class term :361-364, the rectified term
// :140-164, the batch mean :112 and the detail rows :115-138.  The backward is the closed form TF
// autodiff yields for those operators; no gradient flows through IoU or masks.
//
// Shape of the kernel (B200).  profiles/r01_dense_variants.txt lists the shapes tried before this one and what each
// cost (CTA tiles staged behind a barrier; zero-fill followed by partial-sector value stores; logits through L1;
// per-lane owner path; a dedicated zero-fill warp; 64-register / 32-warp build; smaller CTAs).
//  * everything per-image is prepared by the assignment kernel (assign.cu): the valid labels in each head's grid
//    units (gtab), their count, and the object mask as one bit per record (objbits);
//  * the packed tensor is a flat array of records of L floats.  A WARP owns `cpw` consecutive chunks of 32
//    consecutive records (a chunk is 128*L bytes: a sector-aligned span), independent of every other warp -- no CTA
//    barrier in the loop; the logits and object bits of the warp's next chunk are in flight while it works on the
//    current one.  The labels of the (at most two) images its chunks touch sit in the warp's own shared-memory slots;
//  * a lane owns one record: it reads the two aligned 128-bit words holding the record's 5 box/conf logits with
//    L2-only loads (class logits of non-responsible anchors are never read: their gradient is identically 0),
//    decodes, and decides "IoU < iou_thresh for every label" four labels at a time with a division-free test that is
//    exact outside a 2^-19 band below the threshold (the band and everything above it take the reference expression);
//  * gradient of a chunk: L-5 of a record's L floats are zeros.  One asynchronous bulk (TMA) store per chunk, issued
//    by lane 0 one chunk ahead from a block of zeros in shared memory, covers the whole span; the record's lane then
//    writes the one or two 32-byte sectors around its five values WHOLE (values and zeros together), so no sector is
//    ever partially written.  Without the bulk path (unaligned tensors, short records) lanes write the zeros with
//    128-bit streaming stores under a host-computed mask that skips the value sectors;
//  * the few lanes whose object bit is set are served by the whole warp one after the other: owner terms of their
//    labels added in label order by lanes 0..4, the softmax-CE class gradient with the classes spread over the lanes;
//  * loss sums: registers -> warp shuffles -> one partial row per CTA -> the last CTA to finish folds the rows in a
//    fixed order (double) and writes the 18 detail values + their sum.  No float atomics.
#include <stdlib.h>

#include "sweep_parts.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

#ifndef YV3_DENSE_THREADS
#define YV3_DENSE_THREADS 256
#endif
constexpr int kThreads = YV3_DENSE_THREADS;
constexpr int kWarps = kThreads / 32;
#ifndef YV3_DENSE_ZBLK
#define YV3_DENSE_ZBLK 16384          // bytes of the shared-memory zero block one bulk store reads (multiple of 16)
#endif
#ifndef YV3_DENSE_MINB
// resident CTAs per SM the register budget is set for: 24 warps, 80 registers (32 warps = 64 registers spills and
// measured slower, 98 vs 82 us)
#define YV3_DENSE_MINB (768 / YV3_DENSE_THREADS)
#endif
constexpr int kSmallL = 8;            // records shorter than this are composed in shared memory

// value k (0..4) of a record whose first float sits at position `ph` (0..3) of the 8 floats in (a, b)
__device__ __forceinline__ float pick(const float4& a, const float4& b, int ph, int k) {
    const float f[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    float v = f[k];
    v = (ph == 1) ? f[k + 1] : v;
    v = (ph == 2) ? f[k + 2] : v;
    v = (ph == 3) ? f[k + 3] : v;
    return v;
}

// The 12 floats starting at a 32-byte sector boundary that hold a record's five gradient values at float
// position pos (0..7), zeros elsewhere: three conditional shifts (by 1, 2, 4), selects only.
struct Win12 {
    float4 a, b, c;
};
__device__ __forceinline__ Win12 compose(const float (&gv)[5], int pos) {
    const bool s1 = pos & 1, s2 = pos & 2, s4 = pos & 4;
    float x[6];
    x[0] = s1 ? 0.f : gv[0];
#pragma unroll
    for (int i = 1; i < 5; ++i) x[i] = s1 ? gv[i - 1] : gv[i];
    x[5] = s1 ? gv[4] : 0.f;
    float y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float lo = (i < 6) ? x[i] : 0.f;
        const float hi = (i >= 2) ? x[i - 2] : 0.f;
        y[i] = s2 ? hi : lo;
    }
    float z[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) {
        const float lo = (i < 8) ? y[i] : 0.f;
        const float hi = (i >= 4) ? y[i - 4] : 0.f;
        z[i] = s4 ? hi : lo;
    }
    Win12 w;
    w.a = make_float4(z[0], z[1], z[2], z[3]);
    w.b = make_float4(z[4], z[5], z[6], z[7]);
    w.c = make_float4(z[8], z[9], z[10], z[11]);
    return w;
}

struct Logits {
    float4 a, b;      // VEC: the two aligned words around the record; scalar path: a = t0..t3, b.x = t4
};
template <int VEC>
__device__ __forceinline__ Logits load_logits(const float* __restrict__ predicts, long long grec, int L, bool ok) {
    Logits q;
    q.a = make_float4(0.f, 0.f, 0.f, 0.f);
    q.b = q.a;
    if (ok) {
        if (VEC) {
            // L2-only loads: an L1 miss would pull the whole 128-byte line for 20 useful bytes
            const float4* p = reinterpret_cast<const float4*>(predicts) + ((grec * L) >> 2);
            q.a = __ldcg(p);
            q.b = __ldcg(p + 1);
        } else {
            const float* p = predicts + grec * L;
            q.a = make_float4(__ldcg(p), __ldcg(p + 1), __ldcg(p + 2), __ldcg(p + 3));
            q.b.x = __ldcg(p + 4);
        }
    }
    return q;
}

constexpr int kWideL = 12;            // from here on a 32-byte sector holds values of at most one record

// ZW: a ninth warp does nothing but the zero fill of the CTA's chunks (those without an owning record), so that the
// eight computing warps never queue behind the store stream; the sectors it writes and the sectors the record lanes
// write are disjoint, so the two never need to meet.
template <int VEC, int SGT, int ZW>
__global__ void __launch_bounds__(kThreads + 32 * ZW, YV3_DENSE_MINB) loss_dense_kernel(const __grid_constant__ Geo g,
                                                                           const __grid_constant__ DenseArgs a) {
    extern __shared__ __align__(16) float gt_s[];          // SGT: [2][gt_slot_floats(m)] (slot = image & 1), then the zero block
    __shared__ float red_s[kWarps + 1][12];
    __shared__ float stage_s[VEC ? 1 : kWarps * 32 * kSmallL];
    const int nthreads = kThreads + 32 * ZW, nwarps = kWarps + ZW;
    const bool with_grad = a.grad != nullptr && !(a.debug & 8);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int L = g.L, m = a.m;
    const float thr = g.iou_thresh;
    const float thr_lo = __fmul_rn(thr, 0.99999809265136719f);      // thr * (1 - 2^-19), see surely_below
    const bool wide = VEC && L >= kWideL;

    float an0 = 0.f, an1 = 0.f, an2 = 0.f;       // noobj sums per head
    float ar0 = 0.f, ar1 = 0.f, ar2 = 0.f;       // sum of squared raw box logits per head
    float ac0 = 0.f, ac1 = 0.f, ac2 = 0.f;       // class sums per head

    // zero-fill bookkeeping of the wide path: this lane's 128-bit words of a chunk are lane, lane + 32, ...;
    // word c lies in sector c >> 1, which starts at float 8 * (c >> 1) of the chunk = position q of a record.
    // The sector carries box/conf values iff q <= 4 (a record starts at or just before it) or q >= L - 7
    // (the next record starts inside it); those sectors are written whole by the record's lane instead.
    // A chunk starts on a record boundary, so the pattern is the same for every chunk.
    // The pattern is one bit per loop trip, computed on the host (a.zlo / a.zhi); trips beyond 64 (L > 256) are
    // tested on the fly.
    const int dq = 128 % L;
    const int trips = (8 * L + 31) >> 5;             // 128-bit words of a full chunk / 32 lanes
    const unsigned zlo = a.zlo[lane], zhi = a.zhi[lane];
    int q64 = 0;                                     // the position after 64 trips
    if (trips > 64) {
        int q = (8 * (lane >> 1)) % L;
        for (int k = 0; k < 64; ++k) { q += dq; q -= (q >= L) ? L : 0; }
        q64 = q;
    }

    // a warp owns a.cpw consecutive chunks (the launcher sizes the grid so that every warp gets the same number):
    // it stays inside one or two images, whose labels it keeps in its shared-memory slots
    const bool zero_warp = ZW && warp == kWarps;
    int cur = zero_warp ? 0 : (blockIdx.x * kWarps + warp) * a.cpw;
    const int cend = zero_warp ? 0 : min(cur + a.cpw, a.num_chunks);
    int nxt = cur + 1;
    float* slots = gt_s;
    Logits qv = load_logits<VEC>(a.predicts, (long long)cur * 32 + lane, L,
                                 cur < cend && (long long)cur * 32 + lane < a.total_rec);
    unsigned objn = (cur < cend) ? __ldg(a.objbits + cur) : 0u;

    // bulk zero fill (a.bulk): a block of zeros in shared memory is the source of one asynchronous bulk (TMA)
    // store per chunk, issued by lane 0 one chunk ahead; the record lanes later overwrite the sectors around
    // their values whole.  The issuing warp never waits on the store queue.
    const bool bulk = VEC && a.bulk && with_grad && wide;
    float* zblk = gt_s + (SGT ? (size_t)2 * gt_slot_floats(m) : 0);
    const unsigned chunk_bytes = 128u * (unsigned)L;
    if (SGT) {
        // the label slots of the (at most two) images the CTA's records lie in, copied as they are (the assignment
        // kernel wrote them in this layout, count and padding included): one 128-bit load per thread, all in flight
        // together with the logits above
        const int slot_f4 = gt_slot_floats(m) >> 2;
        const long long crec0 = (long long)blockIdx.x * kWarps * a.cpw * 32;
        const long long crec1 = min(crec0 + (long long)kWarps * a.cpw * 32, a.total_rec) - 1;
        const int ia = (int)(crec0 / g.rpi), ib = (int)(crec1 / g.rpi);
        const float4* src = reinterpret_cast<const float4*>(a.gtab);
        float4* dst = reinterpret_cast<float4*>(gt_s);
        for (int i = tid; i < 2 * slot_f4; i += nthreads) {
            const int second = i >= slot_f4 ? 1 : 0;
            const int img = ia + second;
            if (img <= ib) dst[(size_t)(img & 1) * slot_f4 + (i - second * slot_f4)] = __ldg(src + (size_t)img * slot_f4 + (i - second * slot_f4));
        }
    }
    if (bulk) {
        for (int i = tid; i < a.zero_bytes / 4; i += nthreads) zblk[i] = 0.f;
        fence_async_smem();
    }
    if (SGT || bulk) __syncthreads();
    if (bulk) {
        if (lane == 0 && cur < cend && cur != a.num_chunks - 1) {
            char* dst = reinterpret_cast<char*>(a.grad) + (size_t)cur * chunk_bytes;
            for (unsigned o = 0; o < chunk_bytes; o += (unsigned)a.zero_bytes)
                bulk_s2g(dst + o, zblk, min((unsigned)a.zero_bytes, chunk_bytes - o));
            bulk_commit();
        }
    }

    if (zero_warp && with_grad && wide && !(a.debug & 512)) {
        const int c0 = blockIdx.x * kWarps * a.cpw, c1 = min(c0 + kWarps * a.cpw, a.num_chunks);
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int c = c0; c < c1; ++c) {
            // chunks with an owning record (their class gradient is written on top of the zeros) and the tensor's
            // last chunk stay with the computing warp
            if (c == a.num_chunks - 1 || __ldg(a.objbits + c) != 0u) continue;
            float4* gl = reinterpret_cast<float4*>(a.grad) + (size_t)c * 8 * L + lane;
            unsigned mk = zlo;
            const int t32 = min(trips, 32);
#pragma unroll 8
            for (int k = 0; k < t32; ++k) {
                if (mk & 1u) __stcs(gl + 32 * k, z4);
                mk >>= 1;
            }
            if (trips > 32) {
                mk = zhi;
                const int t64 = min(trips, 64);
                for (int k = 32; k < t64; ++k) {
                    if (mk & 1u) __stcs(gl + 32 * k, z4);
                    mk >>= 1;
                }
                int q = q64;
                for (int k = 64; k < trips; ++k) {
                    if (lane + 32 * k < 8 * L && q > 4 && q < L - 7) __stcs(gl + 32 * k, z4);
                    q += dq;
                    q -= (q >= L) ? L : 0;
                }
            }
        }
    }

    while (cur < cend) {
        const int chunk = cur;
        const unsigned objw = objn;                        // rec0 is a multiple of 32: one word per chunk
        const long long rec0 = (long long)chunk * 32;
        const int nrec = (int)min(32LL, a.total_rec - rec0);
        const long long grec = rec0 + lane;
        const bool ok = lane < nrec;
        const long long off = grec * L;                    // first float of the record
        const int ph = VEC ? (int)(off & 3) : 0;
        const float t0 = VEC ? pick(qv.a, qv.b, ph, 0) : qv.a.x, t1 = VEC ? pick(qv.a, qv.b, ph, 1) : qv.a.y;
        const float t2 = VEC ? pick(qv.a, qv.b, ph, 2) : qv.a.z, t3 = VEC ? pick(qv.a, qv.b, ph, 3) : qv.a.w;
        const float t4 = VEC ? pick(qv.a, qv.b, ph, 4) : qv.b.x;
        // the tensor's last chunk may end off a sector boundary: it takes the generic path
        const bool fast = wide && (chunk != a.num_chunks - 1);

        // ---- zeros first: 16 B per store, nothing waits for them ---------------------------------------
        const int nfl = nrec * L;
        if (with_grad && L >= kSmallL) {
            float* gspan = a.grad + (size_t)rec0 * L;
            if (VEC) {
                float4* gdst = reinterpret_cast<float4*>(gspan);
                const int n16 = nfl >> 2;
                const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (fast && bulk) {
                    // zeros already on their way (issued one chunk ahead)
                } else if (fast && ZW && objw == 0u) {
                    // the CTA's zero warp covers this chunk
                } else if (fast) {
                    float4* gl = gdst + lane;
                    unsigned mk = zlo;
                    const int t32 = min(trips, 32);
#pragma unroll 8
                    for (int k = 0; k < t32; ++k) {
                        if (mk & 1u) __stcs(gl + 32 * k, z4);
                        mk >>= 1;
                    }
                    if (trips > 32) {
                        mk = zhi;
                        const int t64 = min(trips, 64);
                        for (int k = 32; k < t64; ++k) {
                            if (mk & 1u) __stcs(gl + 32 * k, z4);
                            mk >>= 1;
                        }
                        int q = q64;
                        for (int k = 64; k < trips; ++k) {
                            if (lane + 32 * k < n16 && q > 4 && q < L - 7) __stcs(gl + 32 * k, z4);
                            q += dq;
                            q -= (q >= L) ? L : 0;
                        }
                    }
                } else {
#pragma unroll 4
                    for (int c = lane; c < n16; c += 32) __stcs(gdst + c, z4);
                    if (lane < (nfl & 3)) gspan[(n16 << 2) + lane] = 0.f;
                }
            } else {
                for (int c = lane; c < nfl; c += 32) __stcs(gspan + c, 0.f);
            }
        }
        // ---- the next chunk's logits go in flight; this chunk's object bits ------------------------------
        qv = load_logits<VEC>(a.predicts, (long long)nxt * 32 + lane, L,
                              nxt < cend && (long long)nxt * 32 + lane < a.total_rec);
        objn = (nxt < cend) ? __ldg(a.objbits + nxt) : 0u;
        const bool ahead = bulk && nxt < cend && nxt != a.num_chunks - 1;
        if (ahead && lane == 0) {
            char* dst = reinterpret_cast<char*>(a.grad) + (size_t)nxt * chunk_bytes;
            for (unsigned o = 0; o < chunk_bytes; o += (unsigned)a.zero_bytes)
                bulk_s2g(dst + o, zblk, min((unsigned)a.zero_bytes, chunk_bytes - o));
            bulk_commit();
        }
        const bool is_obj = ok && ((objw >> lane) & 1u);     // :328 scatter of owned anchors

        // ---- record phase: lane = record (loss_record, shared with the fused sweep) --------------------------
        float gv[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
        // one division per chunk (uniform), lanes step from there
        const int n_first = (int)((unsigned)rec0 / (unsigned)g.rpi);
        const int r_first = (int)((unsigned)rec0 - (unsigned)n_first * (unsigned)g.rpi);
        int n_img = n_first, r_img = r_first + lane, head = 0;
        if (ok) {
            while (r_img >= g.rpi) { r_img -= g.rpi; ++n_img; }
            const RecLoc p = locate(g, r_img);
            head = p.h;
            const BoxDec d = decode_box(g, p, t0, t1, t2, t3, t4);
            const float* slot = SGT ? slots + (n_img & 1) * gt_slot_floats(m) : a.gtab + (size_t)n_img * gt_slot_floats(m);
            const float rc = (p.h == 0) ? a.rect_c[0] : ((p.h == 1) ? a.rect_c[1] : a.rect_c[2]);
            float noobj, sq;
            loss_record<SGT != 0>(g, p, d, t0, t1, t2, t3, slot, m, is_obj, with_grad, a.kappa, rc, a.rect_active, thr, thr_lo,
                                  gv[0], gv[1], gv[2], gv[3], gv[4], noobj, sq);
            an0 += (p.h == 0) ? noobj : 0.f; an1 += (p.h == 1) ? noobj : 0.f; an2 += (p.h == 2) ? noobj : 0.f;
            ar0 += (p.h == 0) ? sq : 0.f; ar1 += (p.h == 1) ? sq : 0.f; ar2 += (p.h == 2) ? sq : 0.f;
        }

        // ---- the record's five gradient values ---------------------------------------------------------
        if (with_grad) {
            if (fast) {
                // whole sectors: disjoint from the lanes' zero fill; after the bulk zero fill of this chunk has
                // completed (lane 0 waits for its bulk group, the warp barrier passes that on)
                if (bulk) {
                    if (lane == 0) {
                        if (ahead) bulk_wait_all<1>(); else bulk_wait_all<0>();
                        asm volatile("fence.proxy.async.global;" ::: "memory");
                    }
                    __syncwarp();
                }
                if (ok && !(a.debug & 256)) {
                    const int pos = (int)(off & 7);
                    const Win12 w = compose(gv, pos);
                    float4* o = reinterpret_cast<float4*>(a.grad) + ((off >> 3) << 1);
                    __stcs(o, w.a);
                    __stcs(o + 1, w.b);
                    if (pos >= 4) {
                        __stcs(o + 2, w.c);
                        __stcs(o + 3, make_float4(0.f, 0.f, 0.f, 0.f));
                    }
                }
            } else if (L >= kSmallL) {
                __syncwarp();                 // every lane's zero stores are ordered before the values
                if (ok) {
                    float* o = a.grad + (size_t)off;
#pragma unroll
                    for (int c = 0; c < 5; ++c) o[c] = gv[c];
                }
            } else {
                // short records (class_num < 3): every float of the span carries a value or sits next to one;
                // compose the chunk in shared memory and write it out coalesced
                float* st = stage_s + warp * 32 * kSmallL;
                if (ok) {
                    float* r = st + lane * L;
                    r[0] = gv[0]; r[1] = gv[1]; r[2] = gv[2]; r[3] = gv[3]; r[4] = gv[4];
                    for (int c = 5; c < L; ++c) r[c] = 0.f;
                }
                __syncwarp();
                float* gspan = a.grad + (size_t)rec0 * L;
                for (int c = lane; c < nfl; c += 32) gspan[c] = st[c];
            }
        }

        // ---- records that own a label (object bit set): the whole warp serves them one after the other ------
        // (owner_record, shared with the fused sweep: owner-side box/conf terms of assign.cu added in label order,
        // duplicates add, rule T3; then the class term :361-364 and its softmax-CE gradient)
        unsigned ob = __ballot_sync(0xffffffffu, is_obj);
        if (ob && !(a.debug & 4)) {
            if (with_grad) __syncwarp();      // the chunk's stores above are ordered before the read-modify-writes
            while (ob) {
                const int src = __ffs(ob) - 1;
                ob &= ob - 1;
                const int o_n = __shfl_sync(0xffffffffu, n_img, src);
                const int o_r = __shfl_sync(0xffffffffu, r_img, src);
                const int o_h = __shfl_sync(0xffffffffu, head, src);
                const size_t o_off = (size_t)(rec0 + src) * L;
                const float gbase = (with_grad && lane < 5) ? __ldcg(a.grad + o_off + lane) : 0.f;
                owner_record(g, a, o_n, o_r, o_h, o_off, gbase, with_grad, lane, ac0, ac1, ac2);
            }
        }
        cur = nxt;
        nxt += 1;
    }

    // ---- reduction: thread -> warp -> CTA partial row (loss_fold_kernel folds the rows) ----------------
    float v[9] = {an0, an1, an2, ac0, ac1, ac2, ar0, ar1, ar2};
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        const float w = warp_sum(v[k]);
        if (lane == 0) red_s[warp][k] = w;
    }
    __syncthreads();
    if (tid < 9) {
        float acc = 0.f;
        for (int w = 0; w < nwarps; ++w) acc += red_s[w][tid];
        a.partials[((size_t)a.row_base + blockIdx.x) * kPartialFloats + tid] = acc;
    }
}

// The fold of the per-CTA partial rows (the launch boundary orders it after every row is written): thread t takes
// rows t, t + T, ... in index order, in double; lanes fold by shuffles, warps through shared memory -- a fixed order,
// hence a deterministic sum -- then the 18 detail values (:115-138) and their sum (:137).
constexpr int kFoldThreads = 256;
constexpr int kFoldWideCtas = 48;           // first level of the two-level fold (tables of tens of thousands of rows)

// per-thread partial sums (double) -> the 18 detail values (:115-138) and their sum (:137): lanes fold by shuffles,
// warps through shared memory, in a fixed order.  part[]: noobj 0..2, cls 3..5, rect 6..8, xy 9..11, wh 12..14, obj 15..17
__device__ __forceinline__ void fold_tail(const Geo& g, const DenseArgs& a, double (&part)[18], int tid) {
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int nwarps = kFoldThreads / 32;
    __shared__ double fold_s[nwarps][18];
    for (int i = tid; i < a.n; i += kFoldThreads) {
#pragma unroll
        for (int k = 0; k < 9; ++k) part[9 + k] += (double)__ldcg(a.img_terms + (size_t)i * 9 + k);   // xy, wh, obj
    }
#pragma unroll
    for (int k = 0; k < 18; ++k) {
        double x = part[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) fold_s[warp][k] = x;
    }
    __syncthreads();
    __shared__ float fin[18];
    if (tid < 18) {
        const int row = tid / 3, h = tid - row * 3;       // rows: xy, wh, noobj, obj, cls, rect
        const int src = (row == 0) ? 9 + h : (row == 1) ? 12 + h : (row == 2) ? h : (row == 3) ? 15 + h
                      : (row == 4) ? 3 + h : 6 + h;
        double acc = 0.0;
        for (int w = 0; w < nwarps; ++w) acc += fold_s[w][src];
        if (row == 2) acc *= (double)g.lw[h][2];
        else if (row == 4) acc *= (double)g.lw[h][4];
        else if (row == 5) acc = a.rect_active ? acc * (double)g.rw[h] : 0.0;
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
    }
}

// rows b0, b0 + stride, ... < b1 of the partial table into part[0..8], in index order, in double
__device__ __forceinline__ void fold_rows(const float* partials, int b0, int b1, int stride, double (&part)[18]) {
#pragma unroll 4
    for (int b = b0; b < b1; b += stride) {
        const float4* row = reinterpret_cast<const float4*>(partials + (size_t)b * kPartialFloats);
        const float4 r0 = __ldcg(row), r1 = __ldcg(row + 1), r2 = __ldcg(row + 2);
        part[0] += (double)r0.x; part[1] += (double)r0.y; part[2] += (double)r0.z;      // noobj
        part[3] += (double)r0.w; part[4] += (double)r1.x; part[5] += (double)r1.y;      // cls
        part[6] += (double)r1.z; part[7] += (double)r1.w; part[8] += (double)r2.x;      // rect
    }
}

// The fold of the per-CTA partial rows (the launch boundary orders it after every row is written): thread t takes
// rows t, t + T, ... in index order, in double -- a fixed order, hence a deterministic sum.
__global__ void __launch_bounds__(kFoldThreads) loss_fold_kernel(const __grid_constant__ Geo g,
                                                                 const __grid_constant__ DenseArgs a, int nrows) {
    double part[18];
#pragma unroll
    for (int k = 0; k < 18; ++k) part[k] = 0.0;
    fold_rows(a.partials, threadIdx.x, nrows, kFoldThreads, part);
    fold_tail(g, a, part, threadIdx.x);
}

// Two levels for long tables (one row per warp of the fused sweep: a single SM would spend 20 us pulling a megabyte
// through its L2 port): CTA c folds its contiguous slice of rows into nine doubles (fixed order inside the CTA), the
// last CTA to finish -- a ticket in status[1], zeroed with the status word by the assignment phase -- adds the slices
// in index order and finishes like the one-level kernel.  The result does not depend on which CTA comes last.
__global__ void __launch_bounds__(kFoldThreads) loss_fold_wide_kernel(const __grid_constant__ Geo g,
                                                                      const __grid_constant__ DenseArgs a, int nrows,
                                                                      double* __restrict__ slices) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int nwarps = kFoldThreads / 32;
    __shared__ double lvl_s[nwarps][9];
    __shared__ int last_s;
    double part[18];
#pragma unroll
    for (int k = 0; k < 18; ++k) part[k] = 0.0;
    const int per = (nrows + gridDim.x - 1) / gridDim.x;
    const int b0 = blockIdx.x * per, b1 = min(b0 + per, nrows);
    fold_rows(a.partials, b0 + tid, b1, kFoldThreads, part);
#pragma unroll
    for (int k = 0; k < 9; ++k) {
        double x = part[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (lane == 0) lvl_s[warp][k] = x;
    }
    __syncthreads();
    if (tid < 9) {
        double acc = 0.0;
        for (int w = 0; w < nwarps; ++w) acc += lvl_s[w][tid];
        slices[(size_t)blockIdx.x * 9 + tid] = acc;
        __threadfence();
    }
    __syncthreads();
    if (tid == 0) last_s = (atomicAdd((int*)a.status + 1, 1) == (int)gridDim.x - 1) ? 1 : 0;
    __syncthreads();
    if (!last_s) return;
    __threadfence();
#pragma unroll
    for (int k = 0; k < 18; ++k) part[k] = 0.0;
    // slices to shared memory in parallel (one load per thread), then thread 0 adds them in index order
    __shared__ double all_s[kFoldWideCtas * 9];
    for (int i = tid; i < (int)gridDim.x * 9; i += kFoldThreads) all_s[i] = __ldcg(slices + i);
    __syncthreads();
    if (tid == 0) {
        for (int c = 0; c < (int)gridDim.x; ++c) {
#pragma unroll
            for (int k = 0; k < 9; ++k) part[k] += all_s[c * 9 + k];
        }
        ((int*)a.status)[1] = 0;                           // ready for a fold that is not preceded by the assignment
    }
    fold_tail(g, a, part, tid);
}

template <int VEC, int SGT, int ZW>
static cudaError_t launch_variant(const Geo& g, DenseArgs& a, int sm_count, bool sweep, bool fold, cudaStream_t stream, int* grid_out) {
    a.bulk = (VEC && !ZW && g.L >= kWideL && !(a.debug & 64)) ? 1 : 0;
    a.zero_bytes = a.bulk ? (int)min(128LL * g.L, (long long)YV3_DENSE_ZBLK) : 0;
    const size_t smem = (SGT ? (size_t)2 * gt_slot_floats(a.m) * sizeof(float) : 0) + (size_t)a.zero_bytes;
    cudaError_t err = cudaFuncSetAttribute(loss_dense_kernel<VEC, SGT, ZW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    (void)sm_count;
    const int grid = (int)max(1LL, ((long long)a.num_chunks + (long long)a.cpw * kWarps - 1) / ((long long)a.cpw * kWarps));
    if (sweep) {
        loss_dense_kernel<VEC, SGT, ZW><<<grid, kThreads + 32 * ZW, smem, stream>>>(g, a);
        err = cudaGetLastError();
        if (err != cudaSuccess) return err;
    }
    if (grid_out) *grid_out = grid;
    if (fold) loss_fold_kernel<<<1, kFoldThreads, 0, stream>>>(g, a, a.row_base + grid);      // the grid is a function of the arguments
    return cudaGetLastError();
}

cudaError_t launch_loss_fold(const Geo& g, const DenseArgs& a, int nrows, cudaStream_t stream) {
    if (nrows > 16 * kFoldThreads && a.fold_slices) loss_fold_wide_kernel<<<kFoldWideCtas, kFoldThreads, 0, stream>>>(g, a, nrows, a.fold_slices);
    else loss_fold_kernel<<<1, kFoldThreads, 0, stream>>>(g, a, nrows);
    return cudaGetLastError();
}

cudaError_t launch_loss_dense(const Geo& g, DenseArgs a, int sm_count, bool sweep, bool fold, cudaStream_t stream, int* grid_out) {
    const int max_grid = a.max_grid > 0 ? a.max_grid : kMaxGrid;
    a.num_chunks = (int)((a.total_rec + 31) / 32);
    // zero-fill pattern of the whole-sector path (see the kernel): bit k of lane l <=> 128-bit word l + 32 k of a
    // chunk lies in a sector without box/conf values
    const int L = g.L;
    for (int lane = 0; lane < 32; ++lane) {
        unsigned lo = 0, hi = 0;
        for (int k = 0; k < 64; ++k) {
            const int c = lane + 32 * k;
            if (c >= 8 * L) break;
            const int q = (8 * (c >> 1)) % L;
            const bool z = q > 4 && q < L - 7;
            if (k < 32) lo |= z ? (1u << k) : 0u; else hi |= z ? (1u << (k - 32)) : 0u;
        }
        a.zlo[lane] = lo;
        a.zhi[lane] = hi;
    }
    // one chunk per warp and as many CTAs as that takes: short-lived CTAs give their slot back as soon as their slowest
    // warp is done, and the hardware scheduler balances the rest (measured best on B200 once the final fold left the
    // kernel: 68 / 70 / 75 us at 1 / 2 / 3 chunks per warp, profiles/r01_dense_variants.txt); more chunks per warp
    // only when the grid would exceed the partial-row table
    // short records (class_num < 11): the per-warp fixed costs weigh more against 32 * L * 4 bytes of payload;
    // measured at L = 5, batch 128 on B200: 85 / 74 / 69 us at 1 / 2 / 4 chunks per warp
    a.cpw = (g.L < 16 && a.num_chunks >= 148 * 64) ? 4 : 1;
    while (((long long)a.num_chunks + (long long)a.cpw * kWarps - 1) / ((long long)a.cpw * kWarps) > max_grid) ++a.cpw;
    if (const char* e = getenv("YV3_CPW")) {                       // developer override
        const int v = atoi(e);
        const long long need = ((long long)a.num_chunks + (long long)v * kWarps - 1) / ((long long)v * kWarps);
        if (v >= 1 && need <= max_grid) a.cpw = v;
    }
    // labels in shared memory when they fit and the records of a CTA lie in at most two images (two slots per CTA)
    const bool sgt = a.m <= 32 && g.rpi >= (long long)kWarps * a.cpw * 32 && !(a.debug & 16);
    const bool vec = a.vec_ok && g.L >= kSmallL;
    // measured on B200 (profiles/r01_dense_variants.txt): the zero warp does not pay (88 vs 83 us at cfg2); kept as a
    // developer switch (YV3_DEBUG bit 128)
    const bool zw = vec && g.L >= kWideL && a.grad != nullptr && (a.debug & 128);
    if (vec && zw) return sgt ? launch_variant<1, 1, 1>(g, a, sm_count, sweep, fold, stream, grid_out) : launch_variant<1, 0, 1>(g, a, sm_count, sweep, fold, stream, grid_out);
    if (vec) return sgt ? launch_variant<1, 1, 0>(g, a, sm_count, sweep, fold, stream, grid_out) : launch_variant<1, 0, 0>(g, a, sm_count, sweep, fold, stream, grid_out);
    return sgt ? launch_variant<0, 1, 0>(g, a, sm_count, sweep, fold, stream, grid_out) : launch_variant<0, 0, 0>(g, a, sm_count, sweep, fold, stream, grid_out);
}

}  // namespace yv3
