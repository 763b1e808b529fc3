// micro-benchmark: how fast can a kernel on GPU 0 READ memory of GPU 1 over NVLink, by access pattern
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); return 1; } } while (0)

template <int MODE>
__global__ void __launch_bounds__(512) k_grid_stride(const uint64_t *__restrict__ p, size_t n, unsigned long long *out) {
    unsigned long long acc = 0;
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, step = (size_t)gridDim.x * blockDim.x;
    if (MODE == 2) {
        const uint4 *q = (const uint4 *)p;
        for (size_t j = i; j < n / 2; j += step) { uint4 v = q[j]; acc += v.x + v.z; }
    } else {
        for (; i + 3 * step < n; i += 4 * step) {
            uint64_t a, b, c, d;
            if (MODE == 1) { a = __ldcs(p + i); b = __ldcs(p + i + step); c = __ldcs(p + i + 2 * step); d = __ldcs(p + i + 3 * step); }
            else { a = p[i]; b = p[i + step]; c = p[i + 2 * step]; d = p[i + 3 * step]; }
            acc += a + b + c + d;
        }
    }
    if (acc == 0x1234567) *out = acc;
}
// merge-like: CTA b reads its own contiguous segment of seg elements, U loads per thread in flight
template <int U, bool PREFETCH>
__global__ void __launch_bounds__(512) k_cta_segment(const uint64_t *__restrict__ p, size_t seg, unsigned long long *out) {
    const uint64_t *s = p + (size_t)blockIdx.x * seg;
    if (PREFETCH) for (size_t o = threadIdx.x * 16; o < seg; o += 512 * 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(s + o));
    unsigned long long acc = 0;
    for (size_t j0 = 0; j0 < seg; j0 += (size_t)U * 512) {
        uint64_t v[U];
#pragma unroll
        for (int u = 0; u < U; u++) { size_t j = j0 + u * 512 + threadIdx.x; v[u] = j < seg ? s[j] : 0; }
#pragma unroll
        for (int u = 0; u < U; u++) acc += v[u];
    }
    if (acc == 0x1234567) *out = acc;
}
int main() {
    int nd = 0; CK(cudaGetDeviceCount(&nd));
    if (nd < 2) { printf("need 2 GPUs\n"); return 0; }
    const size_t bytes = (size_t)512 << 20, n = bytes / 8;
    uint64_t *remote, *local; unsigned long long *out;
    CK(cudaSetDevice(1)); CK(cudaMalloc(&remote, bytes)); CK(cudaMemset(remote, 1, bytes)); CK(cudaDeviceSynchronize());
    CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&local, bytes)); CK(cudaMemset(local, 1, bytes)); CK(cudaMalloc(&out, 8));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto run = [&](const char *name, auto launch) {
        for (int where = 0; where < 2; where++) {
            const uint64_t *p = where ? remote : local;
            launch(p); cudaDeviceSynchronize();
            cudaEventRecord(e0); for (int r = 0; r < 3; r++) launch(p); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
            printf("%-44s %-6s %8.3f ms  %8.1f GB/s\n", name, where ? "remote" : "local", ms, bytes / ms / 1e6);
        }
        cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("  error %s\n", cudaGetErrorString(e));
    };
    run("grid-stride 8B x4, ld", [&](const uint64_t *p) { k_grid_stride<0><<<148 * 4, 512>>>(p, n, out); });
    run("grid-stride 8B x4, ld.cs", [&](const uint64_t *p) { k_grid_stride<1><<<148 * 4, 512>>>(p, n, out); });
    run("grid-stride 16B", [&](const uint64_t *p) { k_grid_stride<2><<<148 * 4, 512>>>(p, n, out); });
    const size_t seg = 8192;   // 64 KB per CTA, like one partition piece
    const int nseg = (int)(n / seg);
    run("CTA segment 64KB, U=1", [&](const uint64_t *p) { k_cta_segment<1, false><<<nseg, 512>>>(p, seg, out); });
    run("CTA segment 64KB, U=4", [&](const uint64_t *p) { k_cta_segment<4, false><<<nseg, 512>>>(p, seg, out); });
    run("CTA segment 64KB, U=8", [&](const uint64_t *p) { k_cta_segment<8, false><<<nseg, 512>>>(p, seg, out); });
    run("CTA segment 64KB, U=4 + prefetch.L2", [&](const uint64_t *p) { k_cta_segment<4, true><<<nseg, 512>>>(p, seg, out); });
    // copy engines, for scale
    for (int dir = 0; dir < 2; dir++) {
        cudaEventRecord(e0); for (int r = 0; r < 3; r++) cudaMemcpyAsync(dir ? (void *)remote : (void *)local, dir ? (void *)local : (void *)remote, bytes, cudaMemcpyDefault); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
        printf("%-44s %-6s %8.3f ms  %8.1f GB/s\n", "cudaMemcpyAsync (copy engine)", dir ? "push" : "pull", ms, bytes / ms / 1e6);
    }
    return 0;
}
