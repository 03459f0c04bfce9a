// yv3_nms_static: yv3_nms without a caller workspace and without any allocation.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  The scratch the NMS kernels need
// (sort keys, sorted boxes, row indices, the spill list of kept boxes, the two segment queues and their counters --
// what yv3_nms carves out of the caller's workspace) lives in ONE file-scope device array of
// YV3_NMS_STATIC_SCRATCH_BYTES (4 MiB).  Every GPU that loads the library holds that array until the process exits.
//
//   * host code cannot use the array's name as an address: it is the host-side stand-in of the symbol.  The device
//     address comes from cudaGetSymbolAddress, asked once per device with that device current (cached below), and that
//     pointer is what the launches get;
//   * all calls on one device share the array and it keeps the last call's contents.  The kernels never rely on them:
//     the only words that must start at zero (the queue counters and the dispatch ticket) are cleared by launch_nms on
//     the call's stream, everything else is written before it is read;
//   * two calls that overlap on the GPU (different streams without an ordering between them, or two host threads on two
//     streams) race on the array.  Calls on one stream, or ordered by events, are safe;
//   * a request that needs more than the array holds is refused with YV3_ERR_STATIC_SCRATCH before anything is launched.
//
// No cudaMalloc / cudaFree / allocating library routine is called: argument checks, cudaGetDevice, (first call per
// device) cudaGetSymbolAddress, then launch_nms -- one cudaMemsetAsync of 12 bytes and the kernel launches.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "yv3_kernels.cuh"

using namespace yv3;

namespace {

constexpr size_t kStaticScratchBytes = YV3_NMS_STATIC_SCRATCH_BYTES;
constexpr int kMaxDevices = 64;

// the scratch itself: device code in this file may name it, host code must go through static_scratch_address()
static __device__ __align__(256) unsigned char g_nms_static_scratch[kStaticScratchBytes];

std::mutex g_addr_mu;
void* g_addr[kMaxDevices] = {nullptr};

// device address of the array on the CURRENT device (asked once per device)
int static_scratch_address(void** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return YV3_ERR_CUDA;
    std::lock_guard<std::mutex> lk(g_addr_mu);
    if (!g_addr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_nms_static_scratch);
        if (e != cudaSuccess || !p) {
            if (getenv("YV3_TRACE")) fprintf(stderr, "[yv3] cudaGetSymbolAddress failed: %s\n", cudaGetErrorString(e));
            return YV3_ERR_CUDA;
        }
        g_addr[dev] = p;
    }
    *out = g_addr[dev];
    return YV3_OK;
}

// the NMS members of the post-process workspace, carved from `base` (the members only the filter uses stay null)
PostWs carve_nms_only(void* base, int n, int capacity) {
    PostWs w;
    memset(&w, 0, sizeof(w));
    size_t off = 0;
    char* p = (char*)base;
    auto take = [&](size_t bytes) {
        char* r = p + off;
        off = align_up(off + bytes, 256);
        return r;
    };
    const size_t rows = (size_t)n * (size_t)capacity;
    w.keys = (unsigned long long*)take((size_t)n * nms_cap2(capacity) * 8);
    w.sbox = (float4*)take(rows * 16);
    w.sidx = (int32_t*)take(rows * 4);
    w.kspill = (float4*)take(rows * 16);
    w.wl = (int2*)take((rows + rows / 256 + 16) * 8);      // medium-segment queue [rows], then the long-segment queue
    w.wl_count = (int32_t*)take(256);
    w.bytes = off;
    return w;
}

}  // namespace

extern "C" {

int yv3_nms_static(const float* cand, const int32_t* counts, int count_cols, int n, int32_t capacity, int num_classes,
                   double nms_thresh, uint8_t* keep, void* stream) {
    if (!cand || !counts || !keep) return YV3_ERR_NULL;
    if (count_cols < 1 || count_cols > 8 || num_classes < 1 || num_classes > 16384 || n < 1) return YV3_ERR_SHAPE;
    if (capacity < 1 || capacity > YV3_MAX_CANDIDATES) return YV3_ERR_CAPACITY;
    if ((uintptr_t)cand & 15u) return YV3_ERR_ALIGN;
    if (carve_nms_only(nullptr, n, capacity).bytes > kStaticScratchBytes) return YV3_ERR_STATIC_SCRATCH;
    void* base = nullptr;
    const int rc = static_scratch_address(&base);
    if (rc != YV3_OK) return rc;
    const PostWs ws = carve_nms_only(base, n, capacity);
    const cudaError_t e = launch_nms(cand, counts, count_cols, n, capacity, num_classes, nms_thresh, keep, ws,
                                     (cudaStream_t)stream);
    if (e != cudaSuccess) {
        if (getenv("YV3_TRACE")) fprintf(stderr, "[yv3] launch_nms (static scratch) failed: %s\n", cudaGetErrorString(e));
        return YV3_ERR_CUDA;
    }
    return YV3_OK;
}

}  // extern "C"
