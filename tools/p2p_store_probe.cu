// SM-issued stores to a peer GPU over NVLink: what a plain copy kernel reaches (the ceiling for k_scatter_segments' remote
// part), against the copy engine.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/p2p_store_probe tools/p2p_store_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int UNROLL>
__global__ void k_copy(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x * UNROLL + threadIdx.x;
    for (; i + (UNROLL - 1) * blockDim.x < n; i += (size_t)gridDim.x * blockDim.x * UNROLL) {
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) v[u] = src[i + u * blockDim.x];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) dst[i + u * blockDim.x] = v[u];
    }
}
__global__ void k_copy_tile(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t ntiles) {   // 8 KB tile per CTA through smem
    __shared__ uint4 s[512];
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        s[threadIdx.x] = src[t * 512 + threadIdx.x];
        s[threadIdx.x + 256] = src[t * 512 + 256 + threadIdx.x];
        __syncthreads();
        dst[t * 512 + threadIdx.x] = s[threadIdx.x];
        dst[t * 512 + 256 + threadIdx.x] = s[threadIdx.x + 256];
        __syncthreads();
    }
}
int main() {
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
    const size_t bytes = 1ull << 30, nv = bytes / 16;
    uint4 *a, *b, *loc;
    CK(cudaSetDevice(1)); CK(cudaMalloc(&b, bytes));
    CK(cudaSetDevice(0)); CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&loc, bytes)); CK(cudaMemset(a, 1, bytes));
    CK(cudaDeviceEnablePeerAccess(1, 0));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto time = [&](const char* name, auto&& f) {
        f(); cudaDeviceSynchronize();
        cudaEventRecord(e0); for (int r = 0; r < 5; r++) f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("%-44s %7.1f GB/s\n", name, 5.0 * bytes / 1e9 / (ms * 1e-3));
        return 0;
    };
    time("copy engine 0 -> 1", [&] { cudaMemcpyAsync(b, a, bytes, cudaMemcpyDefault); });
    time("copy engine 0 -> 1, both 2-byte misaligned", [&] { cudaMemcpyAsync((char*)b + 2, (char*)a + 6, bytes - 16, cudaMemcpyDefault); });
    for (int g : {148, 296, 592, 1184, 2368, 9472}) {
        char nm[96];
        snprintf(nm, sizeof nm, "SM stores 0 -> 1, grid %d x 256, unroll 4", g);
        time(nm, [&] { k_copy<4><<<g, 256>>>(a, b, nv); });
    }
    time("SM stores 0 -> 1, grid 1184, unroll 1", [&] { k_copy<1><<<1184, 256>>>(a, b, nv); });
    time("SM stores 0 -> 1, grid 1184, unroll 8", [&] { k_copy<8><<<1184, 256>>>(a, b, nv); });
    time("SM tile copy through smem 0 -> 1, grid 1184", [&] { k_copy_tile<<<1184, 256>>>(a, b, nv / 512); });
    time("SM tile copy through smem 0 -> 1, grid 131072", [&] { k_copy_tile<<<131072, 256>>>(a, b, nv / 512); });
    time("SM tile copy local, grid 131072", [&] { k_copy_tile<<<131072, 256>>>(a, loc, nv / 512); });
    time("SM loads 1 -> 0 (pull), grid 1184, unroll 4", [&] { k_copy<4><<<1184, 256>>>(b, loc, nv); });
    return 0;
}
