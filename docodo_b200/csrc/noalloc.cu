// dcd_seq_encode_noalloc: the one entry point of the library that allocates nothing (no cudaMalloc / cudaFree / pool block),
// for callers that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.  It has no
// counterpart in the reference.
//
// It is a variant of the K6/K7 encoder (encode.cu, k_encode) for ONE list: ascending 64-bit values -> the 15-bit varint
// difference stream of Builder.Add (IndexSequence.cs:63-84).  The encoder needs scratch for its decoupled look-back (one
// status word per tile of 2048 values) and a tile ticket; here that scratch is a file-scope __device__ array of a fixed size,
// which bounds the list length at DCD_NOALLOC_MAX_VALUES.  Every device that loads the library holds the array until exit.
//
// All calls on a device share the array: the call clears it (stream-ordered memset) before the kernel runs, and two calls that
// overlap on the GPU (different streams, no dependency between them) race on it.  Calls on one stream are safe back to back.
#include "engine.h"
#include "common.cuh"

#include <mutex>

namespace dcd {

constexpr int NA_THREADS = 256;
constexpr int NA_ITEMS = 8;
constexpr int NA_TILE = NA_THREADS * NA_ITEMS;                       // values per tile
constexpr u64 NA_MAX_VALUES = 1ull << 24;                            // == DCD_NOALLOC_MAX_VALUES (include/docodo_b200.h)
constexpr u64 NA_MAX_TILES = NA_MAX_VALUES / NA_TILE;                // 8192
constexpr u64 NA_SCRATCH_WORDS = NA_MAX_TILES + 8;                   // look-back status words + the ticket word: 64 KiB + 64 B

// Host code must not pass this name to a kernel: there it is a host-side stand-in, not a device address.  The address comes
// from cudaGetSymbolAddress (scratch_for_device below), once per device.
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NA_SCRATCH_WORDS];

struct NaSmem {
    u16 stage[NA_TILE * 5 + 8];
    u32 scan[NA_THREADS / 32 + 1];
    u32 tile;
    u32 bad;
    u64 tile_base;
};

DCD_D int na_varint_len(u64 d) {   // ushorts Builder.Add emits for difference d (IndexSequence.cs:68-81)
    int l = 1;
    while (d > 0x7FFF) { d >>= 15; l++; }
    return l;
}

// scratch[0 .. ntiles): look-back status words (zeroed), scratch[NA_MAX_TILES]: tile ticket (zeroed).
// out_n[0] = ushorts of the stream, out_n[1] = 0 ok | 1 out_cap too small | 2 values not ascending (both zeroed by the caller).
__global__ void __launch_bounds__(NA_THREADS, 6) k_encode_list(const u64* __restrict__ values, u64 n, u16* __restrict__ out, u64 out_cap,
                                                              u64* scratch, u64* __restrict__ out_n) {
    extern __shared__ __align__(16) unsigned char smraw[];
    NaSmem& sm = *reinterpret_cast<NaSmem*>(smraw);
    u64* lookback = scratch;
    u32* ticket = reinterpret_cast<u32*>(scratch + NA_MAX_TILES);
    const int tid = threadIdx.x;
    const u32 lane = lane_id();
    if (tid == 0) { sm.tile = atomicAdd(ticket, 1u); sm.bad = 0; }
    __syncthreads();
    const u32 tile = sm.tile;
    const u64 t0 = (u64)tile * NA_TILE;
    const u64 i0 = t0 + (u64)tid * NA_ITEMS;

    u64 v[NA_ITEMS];
#pragma unroll
    for (int j = 0; j < NA_ITEMS; j++) v[j] = i0 + j < n ? __ldg(values + i0 + j) : 0;
    u64 prev = __shfl_up_sync(0xffffffffu, v[NA_ITEMS - 1], 1);
    if (lane == 0) prev = (i0 > 0 && i0 < n) ? __ldg(values + i0 - 1) : 0;   // the list starts from 0 (IndexSequence.cs:66)
    u64 delta[NA_ITEMS];
    u32 lens = 0;   // 4 bits per item: varint length 1..5, 0 = beyond the end
    u32 mine = 0;
    bool bad = false;
#pragma unroll
    for (int j = 0; j < NA_ITEMS; j++) {
        if (i0 + j < n) {
            bad = bad || v[j] < prev;
            delta[j] = v[j] - prev;
            const u32 l = (u32)na_varint_len(delta[j]);
            lens |= l << (4 * j);
            mine += l;
            prev = v[j];
        }
    }
    if (bad) sm.bad = 1;
    u32 tile_total;
    const u32 off0 = block_excl_scan<u32, NA_THREADS>(mine, sm.scan, &tile_total);
    const bool w0 = warp_id() == 0;
    if (w0 && lane == 0) st_volatile_u64(lookback + tile, (tile == 0 ? LB_PREFIX : LB_AGG) | (u64)tile_total);
    {
        u32 off = off0;
#pragma unroll
        for (int j = 0; j < NA_ITEMS; j++) {
            const u32 l = (lens >> (4 * j)) & 7u;
            if (l) {
                u64 d = delta[j];
                for (u32 w = 0; w + 1 < l; w++) { sm.stage[off++] = (u16)(0x8000u | (d & 0x7FFF)); d >>= 15; }
                sm.stage[off++] = (u16)d;
            }
        }
    }
    if (w0) {
        const u64 ex = lookback_warp_wait(lookback, (int)tile, (u64)tile_total);
        if (lane == 0) sm.tile_base = ex;
    }
    __syncthreads();
    const u64 base = sm.tile_base;
    if (tid == 0 && sm.bad) atomicMax(reinterpret_cast<unsigned long long*>(out_n + 1), 2ull);
    if (base + tile_total > out_cap) {
        if (tid == 0) atomicMax(reinterpret_cast<unsigned long long*>(out_n + 1), 1ull);
    } else {
        for (u32 i = tid; i < tile_total; i += NA_THREADS) out[base + i] = sm.stage[i];
    }
    if (t0 + NA_TILE >= n && tid == 0) out_n[0] = base + tile_total;
}

// device address of the scratch array on `device` (which must be current): resolved once per device
static u64* scratch_for_device(int device) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mu;
    static u64* ptrs[MAX_DEVICES] = {};
    if (device < 0 || device >= MAX_DEVICES) throw StatusError(DCD_E_INVALID, "dcd_seq_encode_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            throw StatusError(DCD_E_CUDA, std::string("dcd_seq_encode_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        DCD_CUDA(cudaFuncSetAttribute(k_encode_list, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NaSmem)));
        ptrs[device] = static_cast<u64*>(p);
    }
    return ptrs[device];
}

}  // namespace dcd

using namespace dcd;

extern "C" int32_t dcd_seq_encode_noalloc(dcd_ctx* c, const uint64_t* d_values, uint64_t n, uint16_t* d_out, uint64_t out_cap,
                                          uint64_t* d_out_n) {
    if (!c || !d_out_n || (n && (!d_values || !d_out))) return DCD_E_INVALID;
    try {
        if (n > NA_MAX_VALUES)
            throw StatusError(DCD_E_CAPACITY, "dcd_seq_encode_noalloc: n = " + std::to_string(n) + " exceeds DCD_NOALLOC_MAX_VALUES (" +
                                                  std::to_string(NA_MAX_VALUES) + "): the scratch of this entry point is a fixed device array");
        DCD_CUDA(cudaSetDevice(c->device));
        cudaStream_t s = c->stream;
        u64* scratch = scratch_for_device(c->device);
        DCD_CUDA(cudaMemsetAsync(d_out_n, 0, 16, s));
        if (!n) return DCD_OK;
        const u64 ntiles = (n + NA_TILE - 1) / NA_TILE;
        // the array keeps the previous call's status words and ticket: clear what this call will read
        DCD_CUDA(cudaMemsetAsync(scratch, 0, ntiles * sizeof(u64), s));
        DCD_CUDA(cudaMemsetAsync(scratch + NA_MAX_TILES, 0, 64, s));
        k_encode_list<<<(unsigned)ntiles, NA_THREADS, sizeof(NaSmem), s>>>(reinterpret_cast<const u64*>(d_values), n, d_out, out_cap, scratch,
                                                                          reinterpret_cast<u64*>(d_out_n));
        DCD_CUDA(cudaGetLastError());
        return DCD_OK;
    } catch (const StatusError& e) {
        c->err = e.what();
        return e.code;
    } catch (const CudaError& e) {
        c->err = e.what();
        return DCD_E_CUDA;
    }
}
