// kbench -- stand-alone kernel micro-benchmarks (development tool; not part of the library or the tests).
//   kbench sort  <nkeys> [ncodes]      old 8-bit onesweep vs wide-digit onesweep on Zipf-distributed code ids
// Builds with `make -C docodo_b200/csrc kbench`; run under gpurun.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../docodo_b200/csrc/encode.cuh"
#include "../docodo_b200/csrc/sort.cuh"

using namespace dcd;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ u64 mix64(u64 z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// key i: code = permuted Zipf(1.1)-like rank, coord = 3 * i (ascending)
__global__ void k_gen(u64* keys, u64 n, u32 ncodes) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
        u64 h = mix64(i);
        u32 rank;
        while (true) {
            const double u = ((h >> 11) + 1) * (1.0 / 9007199254740993.0);
            const double r = pow(u, -10.0) - 1.0;
            if (r < (double)ncodes) { rank = (u32)r; break; }
            h = mix64(h);
        }
        const u32 code = (u32)(((u64)rank * 2654435761ull) % ncodes);
        keys[i] = ((u64)code << COORD_BITS) | ((i * 3) & COORD_MASK);
    }
}
__global__ void k_cmp(const u64* a, const u64* b, u64 n, u64* ndiff) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
        if (a[i] != b[i]) atomicAdd(ndiff, 1ull);
}
__global__ void k_check_sorted(const u64* a, u64 n, u64* nbad) {
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += (u64)gridDim.x * blockDim.x)
        if (a[i] > a[i + 1]) atomicAdd(nbad, 1ull);
}

static int bits_for(u64 n) { int b = 1; while ((1ull << b) < n) b++; return b; }

static int bench_sort(u64 n, u32 ncodes) {
    cudaStream_t s;
    CK(cudaStreamCreate(&s));
    u64 *src, *a, *b, *c, *d, *hist, *cnt;
    CK(cudaMalloc(&src, n * 8)); CK(cudaMalloc(&a, n * 8)); CK(cudaMalloc(&b, n * 8)); CK(cudaMalloc(&c, n * 8)); CK(cudaMalloc(&d, n * 8));
    CK(cudaMalloc(&hist, 8 * WIDE_HIST_STRIDE * 8)); CK(cudaMalloc(&cnt, 64));
    k_gen<<<148 * 8, 256, 0, s>>>(src, n, ncodes);
    CK(cudaStreamSynchronize(s));
    const int cb = bits_for(ncodes);
    const SortPlan p8 = make_sort_plan(COORD_BITS, COORD_BITS + cb), pw = make_wide_plan(COORD_BITS, COORD_BITS + cb);
    void *sc8, *scw;
    CK(cudaMalloc(&sc8, sort_scratch_bytes(n))); CK(cudaMalloc(&scw, wide_scratch_bytes(n, pw.npasses)));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int w8 = 0, ww = 0;
    float ms8 = 0, msw = 0, msh = 0;
    for (int it = 0; it < 4; it++) {
        CK(cudaMemcpyAsync(a, src, n * 8, cudaMemcpyDeviceToDevice, s));
        CK(cudaMemsetAsync(hist, 0, 8 * 256 * 8, s));
        launch_histogram(a, n, p8, hist, s);
        cudaEventRecord(e0, s);
        w8 = radix_sort_run(a, b, nullptr, nullptr, n, p8, hist, nullptr, sc8, s, nullptr);
        cudaEventRecord(e1, s);
        CK(cudaStreamSynchronize(s));
        cudaEventElapsedTime(&ms8, e0, e1);
        CK(cudaMemcpyAsync(c, src, n * 8, cudaMemcpyDeviceToDevice, s));
        CK(cudaMemsetAsync(hist, 0, 8 * WIDE_HIST_STRIDE * 8, s));
        cudaEventRecord(e0, s);
        launch_histogram_wide(c, n, pw, hist, s);
        cudaEventRecord(e1, s);
        CK(cudaStreamSynchronize(s));
        cudaEventElapsedTime(&msh, e0, e1);
        cudaEventRecord(e0, s);
        ww = radix_sort_wide(c, d, n, pw, hist, scw, s, nullptr);
        cudaEventRecord(e1, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaGetLastError());
        cudaEventElapsedTime(&msw, e0, e1);
        printf("iter %d: old %d passes %.3f ms (%.0f GB/s) | wide %d passes (%d+%d bits) %.3f ms (%.0f GB/s per pass-equivalent), hist %.3f ms\n", it,
               p8.npasses, ms8, 16.0 * n * p8.npasses / ms8 / 1e6, pw.npasses, pw.bits[0], pw.npasses > 1 ? pw.bits[1] : 0, msw,
               16.0 * n * pw.npasses / msw / 1e6, msh);
    }
    const u64* r8 = w8 ? b : a;
    const u64* rw = ww ? d : c;
    CK(cudaMemsetAsync(cnt, 0, 64, s));
    k_cmp<<<148 * 8, 256, 0, s>>>(r8, rw, n, cnt);
    k_check_sorted<<<148 * 8, 256, 0, s>>>(rw, n, cnt + 1);
    u64 h[2];
    CK(cudaMemcpyAsync(h, cnt, 16, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    printf("sort n=%llu codes=%u: mismatches vs old sort %llu, out-of-order pairs %llu -> %s\n", n, ncodes, h[0], h[1],
           (h[0] == 0 && h[1] == 0) ? "OK" : "FAIL");
    return (h[0] == 0 && h[1] == 0) ? 0 : 1;
}

int main(int argc, char** argv) {
    if (argc < 2) { printf("usage: kbench sort <n> [ncodes]\n"); return 2; }
    if (!strcmp(argv[1], "sort")) {
        const u64 n = argc > 2 ? strtoull(argv[2], 0, 10) : 100000000ull;
        const u32 nc = argc > 3 ? (u32)atoi(argv[3]) : 893287u;
        return bench_sort(n, nc);
    }
    return 2;
}
