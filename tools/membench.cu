// Developer microbenchmark (not part of the library): the memory access pattern of the dense loss pass without its
// arithmetic -- 32-byte reads at a stride of L floats plus a streaming write of the whole tensor -- to find the floor
// the B200 memory system sets for that pattern.    nvcc -O3 -gencode arch=compute_100a,code=sm_100a membench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

// mode bit 0: strided reads, bit 1: zero-fill writes, bit 2: interleaved chunk order (else contiguous per warp),
// bit 3: reads through L1 (__ldg) instead of L2-only, bit 4: plain stores instead of streaming
__global__ void __launch_bounds__(256, 3) pattern(const float4* __restrict__ src, float4* __restrict__ dst, long long nrec,
                                                  int L, int cpw, int mode, float* sink) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * 8 + warp, nw = gridDim.x * 8;
    const int nchunks = (int)((nrec + 31) / 32);
    float acc = 0.f;
    for (int i = 0; i < cpw; ++i) {
        const int chunk = (mode & 4) ? gw + i * nw : gw * cpw + i;
        if (chunk >= nchunks) break;
        const long long rec0 = (long long)chunk * 32;
        if (mode & 2) {
            float4* g = dst + ((rec0 * L) >> 2) + lane;
            const int n16 = 8 * L;
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            if (mode & 16) { for (int c = lane; c < n16; c += 32) g[c - lane] = z; }
            else { for (int c = lane; c < n16; c += 32) __stcs(g + c - lane, z); }
        }
        if (mode & 1) {
            const long long grec = rec0 + lane;
            if (grec < nrec) {
                const float4* p = src + ((grec * L) >> 2);
                float4 a, b;
                if (mode & 8) { a = __ldg(p); b = __ldg(p + 1); } else { a = __ldcg(p); b = __ldcg(p + 1); }
                acc += a.x + b.w;
            }
        }
    }
    if (acc == 123.456f) *sink = acc;
}

// CTA-cooperative variant: a CTA sweeps a contiguous range; all 256 threads write consecutive 4 KB per step
__global__ void __launch_bounds__(256, 3) pattern_cta(const float4* __restrict__ src, float4* __restrict__ dst, long long nrec,
                                                      int L, int rec_per_cta, int mode, float* sink) {
    const long long rec0 = (long long)blockIdx.x * rec_per_cta;
    const long long rec1 = min(nrec, rec0 + rec_per_cta);
    float acc = 0.f;
    if (mode & 2) {
        float4* g = dst + ((rec0 * L) >> 2);
        const long long n16 = ((rec1 - rec0) * L) >> 2;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (long long c = threadIdx.x; c < n16; c += 256) __stcs(g + c, z);
    }
    if (mode & 1) {
        for (long long r = rec0 + threadIdx.x; r < rec1; r += 256) {
            const float4* p = src + ((r * L) >> 2);
            float4 a = __ldcg(p), b = __ldcg(p + 1);
            acc += a.x + b.w;
        }
    }
    if (acc == 123.456f) *sink = acc;
}

int main() {
    const int L = 85;
    const long long nrec = 64LL * 10647;
    const size_t bytes = (size_t)nrec * L * 4;
    float4 *src, *dst; float* sink;
    CK(cudaMalloc(&src, bytes + 64)); CK(cudaMalloc(&dst, bytes + 64)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(src, 0, bytes)); CK(cudaMemset(dst, 0, bytes));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int nchunks = (int)((nrec + 31) / 32);
    auto run = [&](const char* name, int mode, int grid, bool cta, int per) {
        float best = 1e9f, sum = 0.f;
        for (int it = 0; it < 25; ++it) {
            cudaEventRecord(e0);
            if (cta) pattern_cta<<<grid, 256>>>(src, dst, nrec, L, per, mode, sink);
            else pattern<<<grid, 256>>>(src, dst, nrec, L, per, mode, sink);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (it >= 5) { best = ms < best ? ms : best; sum += ms; }
        }
        printf("%-44s grid %5d  best %7.1f us  mean %7.1f us\n", name, grid, best * 1e3f, sum / 20 * 1e3f);
    };
    for (int per_sm = 3; per_sm <= 8; per_sm += (per_sm == 3 ? 1 : 2)) {
        const int slots = 148 * per_sm * 8;
        const int cpw = (nchunks + slots - 1) / slots;
        const int grid = (nchunks + cpw * 8 - 1) / (cpw * 8);
        printf("-- %d CTAs/SM, %d chunks per warp\n", per_sm, cpw);
        run("warp-contiguous  reads only", 1, grid, false, cpw);
        run("warp-contiguous  writes only", 2, grid, false, cpw);
        run("warp-contiguous  reads+writes", 3, grid, false, cpw);
        run("warp-interleaved reads+writes", 7, grid, false, cpw);
        run("warp-interleaved reads(L1)+writes", 15, grid, false, cpw);
        run("warp-interleaved reads+plain writes", 23, grid, false, cpw);
    }
    for (int rpc = 256; rpc <= 4096; rpc *= 2) {
        const int grid = (int)((nrec + rpc - 1) / rpc);
        char nm[64]; snprintf(nm, 64, "cta-sweep %d rec/CTA reads+writes", rpc);
        run(nm, 3, grid, true, rpc);
        snprintf(nm, 64, "cta-sweep %d rec/CTA writes only", rpc);
        run(nm, 2, grid, true, rpc);
    }
    {
        float best = 1e9f;
        for (int it = 0; it < 20; ++it) {
            cudaEventRecord(e0); cudaMemsetAsync(dst, 0, bytes); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best;
        }
        printf("cudaMemset of the tensor: %.1f us\n", best * 1e3f);
    }
    return 0;
}
