// YOLOv3Decoder.decode and LabelDecoder.decode as materialising kernels.
//
// Follows yolov3/yolov3_decoder.py: _unpack :89-117 (pure index math here: the packed tensor is a flat
// stream of records), _decode_xy :143-156, _decode_wh :159-169, _decode_score :172-180,
// _decode_class_prob :183-192 (max-subtracted softmax = exp * (1 / sum), then clip), boxes :138-139,
// and the raw[..., 0:4] side output :84-86.  Labels: yolov3/label_decoder.py:26-60.
//
// Bound: HBM.  4*A*L bytes in, 4*A*(L+8) bytes out per image.  A CTA stages a tile of TR consecutive
// records (a 16-B aligned span of the packed stream) in shared memory with one bulk (TMA) copy, decodes
// it in place -- one thread per record, box part and class softmax -- and streams
// it to the three per-head tensors with coalesced stores.  Several CTAs are resident per SM, so the
// load of one tile overlaps the arithmetic and stores of another.
#include "tile_ops.cuh"
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kDecThreads = 256;

using DecodeLaunch = TileLaunch;

__global__ void __launch_bounds__(kDecThreads) decode_kernel(const __grid_constant__ Geo g, const DecodeArgs a,
                                                             const DecodeLaunch q) {
    extern __shared__ __align__(128) unsigned char dsm[];
    const int L = g.L, C = g.C, tr = q.tr;
    float* tile = (float*)dsm;                                         // [tr][L]
    long long* dst = (long long*)(dsm + align_up((size_t)tr * L * 4, 16));   // [tr] (head<<56 | record in head tensor)
    __shared__ uint64_t bar;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long rec0 = (long long)blockIdx.x * tr;
    const int nrec = (int)min((long long)tr, q.total_rec - rec0);
    const int nfl = nrec * L;
    const float* src = a.predicts + (size_t)rec0 * L;

    tile_load<kDecThreads>(tile, src, nfl, q.tma, &bar);

    // ---- one thread per record: box part, then the class probabilities in place ---------------------
    for (int ri = tid; ri < nrec; ri += kDecThreads) {
        const long long grec = rec0 + ri;
        const int n_img = (int)(grec / g.rpi);
        const int r_img = (int)(grec - (long long)n_img * g.rpi);
        const RecLoc p = locate(g, r_img);
        float* r = tile + (size_t)ri * L;
        const float t0 = r[0], t1 = r[1], t2 = r[2], t3 = r[3], t4 = r[4];
        const BoxDec d = decode_box(g, p, t0, t1, t2, t3, t4);
        const long long rih = (long long)n_img * g.head_rec[p.h] + ((long long)p.y * g.W[p.h] + p.x) * g.B[p.h] + p.b;
        dst[ri] = ((long long)p.h << 56) | rih;
        r[0] = d.x; r[1] = d.y; r[2] = d.w; r[3] = d.h; r[4] = d.conf;
        if (a.raw4[p.h]) reinterpret_cast<float4*>(a.raw4[p.h])[rih] = make_float4(t0, t1, t2, t3);
        if (a.boxes[p.h]) reinterpret_cast<float4*>(a.boxes[p.h])[rih] = make_float4(d.x0, d.y0, d.x1, d.y1);
        if (C > 0) class_softmax_in_place(r + 5, C, g.lo, g.hi);         // decoder :189-191
    }
    __syncthreads();

    // ---- stream the decoded records to the per-head tensors ---------------------------------------
    if (L >= 32) {
        for (int ri = warp; ri < nrec; ri += kDecThreads / 32) {
            const long long d = dst[ri];
            const int h = (int)(d >> 56);
            float* o = a.decoded[h];
            if (!o) continue;
            o += (size_t)(d & 0x00ffffffffffffffLL) * L;
            const float* r = tile + (size_t)ri * L;
            for (int k = lane; k < L; k += 32) __stcs(o + k, r[k]);
        }
    } else {
        for (int i = tid; i < nfl; i += kDecThreads) {
            const int ri = i / L, k = i - ri * L;
            const long long d = dst[ri];
            float* o = a.decoded[(int)(d >> 56)];
            if (o) __stcs(o + (size_t)(d & 0x00ffffffffffffffLL) * L + k, tile[i]);
        }
    }
}

cudaError_t launch_decode(const Geo& g, const DecodeArgs& a, int sm_count, cudaStream_t stream) {
    (void)sm_count;
    DecodeLaunch q;
    q.total_rec = (long long)a.n * g.rpi;
    long long tr = (long long)(65536 / ((size_t)g.L * 4));
    tr = min(tr, (long long)kDecThreads * 4);
    tr &= ~3LL;
    if (tr < 4) {
        // records longer than 16 KB: one record group of 4 per tile as long as it fits the SM
        tr = 4;
        if ((size_t)tr * g.L * 4 > 200 * 1024) return cudaErrorInvalidValue;
    }
    q.tr = (int)tr;
    q.num_tiles = (int)((q.total_rec + tr - 1) / tr);
    q.tma = (((uintptr_t)a.predicts & 15u) == 0) ? 1 : 0;
    const size_t smem = align_up((size_t)tr * g.L * 4, 16) + (size_t)tr * 8 + 128;
    cudaError_t err = cudaFuncSetAttribute(decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    decode_kernel<<<q.num_tiles, kDecThreads, smem, stream>>>(g, a, q);
    return cudaGetLastError();
}

// ---- labels (label_decoder.py:45-60) -------------------------------------------------------------
__global__ void label_decode_kernel(const __grid_constant__ Geo g, const float* __restrict__ targets, long long slots,
                                    float* o0, float* b0, float* o1, float* b1, float* o2, float* b2) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= slots) return;
    const float* t = targets + i * 5;
    const float xn = __ldg(t), yn = __ldg(t + 1), wn = __ldg(t + 2), hn = __ldg(t + 3), cf = __ldg(t + 4);
    float* outs[3] = {o0, o1, o2};
    float* bxs[3] = {b0, b1, b2};
#pragma unroll
    for (int h = 0; h < 3; ++h) {
        const GtDec T = decode_label(g, h, xn, yn, wn, hn);
        if (outs[h]) {
            float* o = outs[h] + i * 5;
            o[0] = T.tx; o[1] = T.ty; o[2] = T.tw; o[3] = T.th; o[4] = cf;
        }
        if (bxs[h]) reinterpret_cast<float4*>(bxs[h])[i] = make_float4(T.box.x0, T.box.y0, T.box.x1, T.box.y1);
    }
}

cudaError_t launch_label_decode(const Geo& g, const float* targets, int n, int m, float* const out[6],
                                cudaStream_t stream) {
    const long long slots = (long long)n * m;
    const int threads = 128;
    const int blocks = (int)((slots + threads - 1) / threads);
    label_decode_kernel<<<blocks, threads, 0, stream>>>(g, targets, slots, out[0], out[1], out[2], out[3], out[4],
                                                         out[5]);
    return cudaGetLastError();
}

// ---- label ingestion (dataset/file_util.py:47-52, :97) ----------------------------------------------------
__global__ void label_batch_kernel(const float* __restrict__ labels, const int32_t* __restrict__ row_offsets,
                                   const int32_t* __restrict__ image_hw, int n, int m, float fh, float fw,
                                   float* __restrict__ targets) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n * m) return;
    const int img = (int)(i / m), slot = (int)(i - (long long)img * m);
    float* o = targets + i * 5;
    const int r0 = row_offsets[img], r1 = row_offsets[img + 1];
    if (slot >= r1 - r0) {                                            // padded_batch(padding_values=-1.)
        o[0] = -1.f; o[1] = -1.f; o[2] = -1.f; o[3] = -1.f; o[4] = -1.f;
        return;
    }
    // src_size_hw = shape[0:2] / image_size;  resize_wh_ratio = src_size_hw[::-1] / reduce_max(src_size_hw)
    const float sh = __fdiv_rn((float)image_hw[2 * img], fh), sw = __fdiv_rn((float)image_hw[2 * img + 1], fw);
    const float mx = fmaxf(sh, sw);
    const float rw = __fdiv_rn(sw, mx), rh = __fdiv_rn(sh, mx);
    const float* t = labels + (size_t)(r0 + slot) * 5;
    // label_xy * ratio + (1 - ratio) / 2.0 ;  label_wh * ratio ;  class unchanged
    o[0] = __fadd_rn(__fmul_rn(__ldg(t + 0), rw), __fmul_rn(__fsub_rn(1.0f, rw), 0.5f));
    o[1] = __fadd_rn(__fmul_rn(__ldg(t + 1), rh), __fmul_rn(__fsub_rn(1.0f, rh), 0.5f));
    o[2] = __fmul_rn(__ldg(t + 2), rw);
    o[3] = __fmul_rn(__ldg(t + 3), rh);
    o[4] = __ldg(t + 4);
}

cudaError_t launch_label_batch(const float* labels, const int32_t* row_offsets, const int32_t* image_hw, int n, int m,
                               int target_h, int target_w, float* targets, cudaStream_t stream) {
    const long long slots = (long long)n * m;
    const int threads = 128;
    label_batch_kernel<<<(int)((slots + threads - 1) / threads), threads, 0, stream>>>(
        labels, row_offsets, image_hw, n, m, (float)target_h, (float)target_w, targets);
    return cudaGetLastError();
}

}  // namespace yv3
