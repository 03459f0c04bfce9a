// All-reduce (sum) of a rank's loss rows over the GPUs of one NVLink domain, as ONE kernel over peer-mapped memory.
//
// The head path has exactly one exchange step: the batch mean of yolov3_loss.py:112 runs over the WHOLE batch, so
// the 20 loss floats of a step (18 detail values, their sum, a pad) must be summed over the ranks that hold the
// shards.  80 bytes: the cost is latency, not bandwidth.  A library collective costs a launch on its own stream plus
// its protocol; worse, round 1 could not capture it in the CUDA graph that holds the step, so the reduction ran once
// per 16-step replay, outside the graph.  This kernel is an ordinary launch -- it sits in the captured step, on the
// side stream next to the fold that produces the row -- and moves the data itself:
//
//   push   every rank stores its row into slot [parity][rank] of EVERY peer's mailbox (plain stores through the
//          peer mapping: NVLink writes are fire-and-forget), fences system-wide, then raises flag [parity][rank] in
//          every peer's mailbox to the call's epoch (st.release.sys);
//   wait   it polls its OWN flags (local memory, ld.acquire.sys) until all ranks' flags show the epoch;
//   sum    it adds the `world` rows of its own mailbox in rank order -- the same order on every rank, so all ranks
//          hold bit-identical sums -- and writes the result.
//
// Two parities: a rank can start call e + 1 (other parity) while a peer still reads call e's slots, and it cannot
// start e + 2 before every peer has raised its flag for e + 1, which a peer does only after it has finished reading e.
// The epoch lives in the mailbox (device memory) and is advanced by the kernel, so a captured graph can be replayed.
// The poll is bounded: a peer that never arrives sets a status bit instead of hanging the GPU.
#include "yv3_kernels.cuh"

namespace yv3 {

constexpr int kPeerMaxWorld = 16;
constexpr int kPeerThreads = 256;

struct PeerTable {
    unsigned char* base[kPeerMaxWorld];     // mailbox of every rank, as mapped into THIS process
};

// mailbox layout (identical on every rank): [epoch u32, status u32, pad to 128] [flags: 2 x world x 32 B] [data: 2 x world x cap floats]
__host__ __device__ inline size_t peer_flags_off() { return 128; }
__host__ __device__ inline size_t peer_data_off(int world) { return 128 + (size_t)2 * world * 32; }
__host__ __device__ inline size_t peer_mailbox_bytes(int world, int cap) {
    return align_up(peer_data_off(world) + (size_t)2 * world * cap * sizeof(float), 256);
}

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(kPeerThreads) peer_allreduce_kernel(const PeerTable t, int world, int rank, int cap,
                                                                      const float* __restrict__ local, int count,
                                                                      float* __restrict__ out) {
    unsigned char* mine = t.base[rank];
    unsigned* epoch_p = reinterpret_cast<unsigned*>(mine);
    unsigned* status_p = epoch_p + 1;
    const unsigned e = *epoch_p + 1u;
    const int parity = (int)(e & 1u);
    const int tid = threadIdx.x;
    // ---- push: my row into slot [parity][rank] of every mailbox
    for (int p = 0; p < world; ++p) {
        float* dst = reinterpret_cast<float*>(t.base[p] + peer_data_off(world)) + ((size_t)parity * world + rank) * cap;
        for (int i = tid; i < count; i += kPeerThreads) dst[i] = local[i];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < world) {
        unsigned* flag = reinterpret_cast<unsigned*>(t.base[tid] + peer_flags_off() + ((size_t)parity * world + rank) * 32);
        st_release_sys(flag, e);
    }
    // ---- wait: every rank's flag in MY mailbox shows this epoch (bounded: ~2 s)
    if (tid < world) {
        const unsigned* flag = reinterpret_cast<const unsigned*>(mine + peer_flags_off() + ((size_t)parity * world + tid) * 32);
        const long long t0 = clock64();
        bool ok = false;
        while (!ok) {
            const unsigned v = ld_acquire_sys(flag);
            ok = (int)(v - e) >= 0;
            if (!ok && clock64() - t0 > 4000000000LL) break;
        }
        if (!ok) atomicOr(status_p, 1u);
    }
    __syncthreads();
    // ---- sum in rank order
    const volatile float* data = reinterpret_cast<const volatile float*>(mine + peer_data_off(world)) + (size_t)parity * world * cap;
    for (int i = tid; i < count; i += kPeerThreads) {
        float acc = 0.f;
        for (int q = 0; q < world; ++q) acc += data[(size_t)q * cap + i];
        out[i] = acc;
    }
    __syncthreads();
    if (tid == 0) *epoch_p = e;
}

}  // namespace yv3

using namespace yv3;

extern "C" {

// bytes of the (symmetric, peer-mapped) mailbox every rank must provide, zero-initialised, for rows of up to `cap` floats
int yv3_peer_mailbox_bytes(int world, int cap, size_t* bytes) {
    if (!bytes) return YV3_ERR_NULL;
    if (world < 1 || world > kPeerMaxWorld || cap < 1) return YV3_ERR_SHAPE;
    *bytes = peer_mailbox_bytes(world, cap);
    return YV3_OK;
}

// out[0..count) = sum over ranks of local[0..count).  mailboxes[r] = rank r's mailbox as mapped into this process.
// Stream-ordered; every rank must make the same sequence of calls.  local may equal out.
int yv3_peer_allreduce(void* const* mailboxes, int world, int rank, int cap, const float* local, int count, float* out,
                       void* stream) {
    if (!mailboxes || !local || !out) return YV3_ERR_NULL;
    if (world < 1 || world > kPeerMaxWorld || rank < 0 || rank >= world || count < 1 || count > cap) return YV3_ERR_SHAPE;
    PeerTable t;
    for (int r = 0; r < kPeerMaxWorld; ++r) t.base[r] = (r < world) ? (unsigned char*)mailboxes[r] : nullptr;
    for (int r = 0; r < world; ++r)
        if (!t.base[r] || ((uintptr_t)t.base[r] & 127u)) return YV3_ERR_ALIGN;
    peer_allreduce_kernel<<<1, kPeerThreads, 0, (cudaStream_t)stream>>>(t, world, rank, cap, local, count, out);
    return cudaGetLastError() == cudaSuccess ? YV3_OK : YV3_ERR_CUDA;
}

// 0, or 1 when a peer did not arrive within the poll bound in some call since the mailbox was zeroed (synchronises)
int yv3_peer_status(const void* mailbox, void* stream) {
    if (!mailbox) return YV3_ERR_NULL;
    unsigned v[2] = {0, 0};
    if (cudaMemcpyAsync(v, mailbox, sizeof(v), cudaMemcpyDeviceToHost, (cudaStream_t)stream) != cudaSuccess) return YV3_ERR_CUDA;
    if (cudaStreamSynchronize((cudaStream_t)stream) != cudaSuccess) return YV3_ERR_CUDA;
    return (int)(v[1] & 1u);
}

}  // extern "C"
