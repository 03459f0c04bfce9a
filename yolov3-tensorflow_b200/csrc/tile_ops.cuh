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

// Class probabilities of one record, in place, by one warp: z -> clip(exp(z - max z) * (1 / sum))
// (yolov3_decoder.py:189-191; tf.nn.softmax's CPU functor multiplies by the reciprocal of the sum).
__device__ __forceinline__ void softmax_clip_inplace(float* z, int C, int lane, float lo, float hi) {
    float mx = -INFINITY;
    for (int k = lane; k < C; k += 32) mx = fmaxf(mx, z[k]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = lane; k < C; k += 32) {
        const float e = expf(__fsub_rn(z[k], mx));
        z[k] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    const float inv = __fdiv_rn(1.0f, sum);
    for (int k = lane; k < C; k += 32) z[k] = clip_ref(__fmul_rn(z[k], inv), lo, hi);
}

// max and FIRST arg-max of q[0..C) by one warp (np.max / np.argmax, yolov3_post_process.py:57-58)
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
