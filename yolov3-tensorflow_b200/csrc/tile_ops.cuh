// Tile staging shared by the decode and the score kernels: a tile is TR consecutive records of the
// packed stream, i.e. one 16-B aligned contiguous span, landed in shared memory with a 1-D bulk (TMA)
// copy when the tensor base is 16-B aligned (coalesced scalar loads otherwise).
#pragma once
#include "yv3_common.cuh"

namespace yv3 {

struct TileLaunch {
    int tr;               // records per tile (multiple of 4)
    int num_tiles;
    long long total_rec;
    int tma;
};

// All threads of the CTA call this; returns after the tile is complete in `tile` (a barrier inside).
template <int THREADS>
__device__ __forceinline__ void tile_load(float* tile, const float* __restrict__ src, int nfl, int tma, uint64_t* bar) {
    const int tid = threadIdx.x;
    if (tma) {
        const unsigned bulk = (unsigned)(((size_t)nfl * 4) & ~(size_t)15);
        if (tid == 0) {
            mbar_init(bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            if (bulk) {
                mbar_expect_tx(bar, bulk);
                bulk_g2s(tile, src, bulk, bar);
            }
        }
        for (int i = (int)(bulk / 4) + tid; i < nfl; i += THREADS) tile[i] = __ldg(src + i);
        __syncthreads();                               // barrier init visible to every waiter
        if (bulk) mbar_wait(bar, 0);
    } else {
        for (int i = tid; i < nfl; i += THREADS) tile[i] = __ldg(src + i);
    }
    __syncthreads();
}

// ---- class probabilities of one record, by ONE THREAD reading its record from shared memory ---------
// (yolov3_decoder.py:189-191).  tf.nn.softmax's CPU functor computes exp(z - max z) * (1 / sum); the sum
// here runs over four interleaved partial sums, (s0 + s1) + (s2 + s3), the same in every kernel that
// needs class probabilities so that the fused filter and the materialised decode agree bit for bit.
// A record is L floats apart from its neighbour's, so for odd L (85) the 32 lanes of a warp hit 32
// different banks.
struct ClassStats {
    float mx;       // max logit
    float inv;      // 1 / sum exp(z - mx)
    int arg;        // first index of the max logit
    int near;       // logits within 2.4e-7 of the max (including it): > 1 means a rounding tie is possible
};

__device__ __forceinline__ ClassStats class_stats(const float* __restrict__ z, int C) {
    ClassStats s;
    float mx = z[0];
    int arg = 0;
    for (int k = 1; k < C; ++k) {
        const float v = z[k];
        if (v > mx) { mx = v; arg = k; }
    }
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int near = 0;
    int k = 0;
    for (; k + 4 <= C; k += 4) {
        const float d0 = __fsub_rn(z[k], mx), d1 = __fsub_rn(z[k + 1], mx);
        const float d2 = __fsub_rn(z[k + 2], mx), d3 = __fsub_rn(z[k + 3], mx);
        near += (d0 >= -2.4e-7f) + (d1 >= -2.4e-7f) + (d2 >= -2.4e-7f) + (d3 >= -2.4e-7f);
        s0 += expf(d0); s1 += expf(d1); s2 += expf(d2); s3 += expf(d3);
    }
    for (; k < C; ++k) {
        const float d = __fsub_rn(z[k], mx);
        near += (d >= -2.4e-7f);
        const float e = expf(d);
        const int m = k & 3;
        if (m == 0) s0 += e; else if (m == 1) s1 += e; else if (m == 2) s2 += e; else s3 += e;
    }
    const float sum = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
    s.mx = mx;
    s.inv = __fdiv_rn(1.0f, sum);
    s.arg = arg;
    s.near = near;
    return s;
}

// The class probabilities of one record IN PLACE (decoder :189-191): z[k] <- clip(exp(z[k] - max) * (1 / sum)).  The same
// values as class_stats + class_prob element by element -- same maximum, the same four interleaved partial sums in the same
// order, the same exp(z - max) -- but every exponential is taken once: it is kept in z[k] between the sum and the scaling.
__device__ __forceinline__ void class_softmax_in_place(float* __restrict__ z, int C, float lo, float hi) {
    float mx = z[0];
    for (int k = 1; k < C; ++k) {
        const float v = z[k];
        if (v > mx) mx = v;                             // (the comparison of class_stats: a NaN never displaces the maximum)
    }
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int k = 0;
    for (; k + 4 <= C; k += 4) {
        const float e0 = expf(__fsub_rn(z[k], mx)), e1 = expf(__fsub_rn(z[k + 1], mx));
        const float e2 = expf(__fsub_rn(z[k + 2], mx)), e3 = expf(__fsub_rn(z[k + 3], mx));
        s0 += e0; s1 += e1; s2 += e2; s3 += e3;
        z[k] = e0; z[k + 1] = e1; z[k + 2] = e2; z[k + 3] = e3;
    }
    for (; k < C; ++k) {
        const float e = expf(__fsub_rn(z[k], mx));
        const int m = k & 3;
        if (m == 0) s0 += e; else if (m == 1) s1 += e; else if (m == 2) s2 += e; else s3 += e;
        z[k] = e;
    }
    const float inv = __fdiv_rn(1.0f, __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3)));
    for (k = 0; k < C; ++k) z[k] = clip_ref(__fmul_rn(z[k], inv), lo, hi);
}

// The same statistics by a whole WARP (all lanes call it, all lanes get the result), bit for bit: class k lives on
// lane k & 31 in slot k >> 5; the four interleaved partial sums of class_stats are rebuilt by lanes 0..3, lane r
// adding the terms of classes r, r + 4, r + 8, ... in increasing order (a shuffle and an add per term); a class past
// the end contributes +0.0, which leaves a positive sum unchanged.
__device__ __forceinline__ ClassStats class_stats_warp(const float* __restrict__ z, int C, int lane) {
    ClassStats s;
    float mx = -INFINITY;
    for (int k = lane; k < C; k += 32) mx = fmaxf(mx, z[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    int arg = 0x7fffffff, near = 0;
    float part = 0.f;                                  // lanes 0..3: partial sum r = lane
    for (int base = 0; base < C; base += 32) {
        const int k = base + lane;
        float e = 0.f;
        if (k < C) {
            const float v = z[k];
            if (v == mx && arg == 0x7fffffff) arg = k;
            const float d = __fsub_rn(v, mx);
            near += (d >= -2.4e-7f) ? 1 : 0;
            e = expf(d);
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const float v = __shfl_sync(0xffffffffu, e, (lane & 3) + 4 * t);
            part += v;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        arg = min(arg, __shfl_xor_sync(0xffffffffu, arg, o));
        near += __shfl_xor_sync(0xffffffffu, near, o);
    }
    const float s0 = __shfl_sync(0xffffffffu, part, 0), s1 = __shfl_sync(0xffffffffu, part, 1);
    const float s2 = __shfl_sync(0xffffffffu, part, 2), s3 = __shfl_sync(0xffffffffu, part, 3);
    const float sum = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
    s.mx = mx;
    s.inv = __fdiv_rn(1.0f, sum);
    s.arg = arg;
    s.near = near;
    return s;
}

// clip(exp(z - mx) * inv): one class probability
__device__ __forceinline__ float class_prob(float z, const ClassStats& s, float lo, float hi) {
    return clip_ref(__fmul_rn(expf(__fsub_rn(z, s.mx)), s.inv), lo, hi);
}

// np.max / np.argmax over the clipped probabilities (yolov3_post_process.py:57-58) without materialising
// them.  The max logit has exp(0) = 1 and multiplying by inv is monotone, so max prob = clip(1 * inv).
// The arg-max is the FIRST index whose rounded probability equals that value: the first max logit, unless
// a logit within rounding distance of the max sits at a lower index and rounds to the same probability.
__device__ __forceinline__ void class_best(const float* __restrict__ z, int C, const ClassStats& s, float lo, float hi,
                                           float& best, int& arg) {
    best = clip_ref(s.inv, lo, hi);
    arg = s.arg;
    if (s.near > 1) {
        for (int k = 0; k < s.arg; ++k) {
            if (class_prob(z[k], s, lo, hi) == best) { arg = k; break; }
        }
    }
}

// max and FIRST arg-max of q[0..C) by one warp (np.max / np.argmax on materialised probabilities)
__device__ __forceinline__ void warp_max_argmax(const float* q, int C, int lane, float& best, int& arg) {
    float v = -INFINITY;
    int a = 0x7fffffff;
    for (int k = lane; k < C; k += 32) {
        const float x = q[k];
        if (x > v || a == 0x7fffffff) { v = x; a = k; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oa = __shfl_xor_sync(0xffffffffu, a, o);
        if (ov > v || (ov == v && oa < a)) { v = ov; a = oa; }
    }
    best = v;
    arg = a;
}

}  // namespace yv3
