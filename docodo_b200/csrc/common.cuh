// docodo_b200 -- shared device/host helpers (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <stdexcept>
#include <string>

typedef unsigned long long u64;
typedef long long i64;
typedef uint32_t u32;
typedef uint16_t u16;
typedef uint8_t u8;

#if defined(__CUDACC__)
#define DCD_HD __host__ __device__ __forceinline__
#define DCD_D __device__ __forceinline__
#else
#define DCD_HD inline
#define DCD_D inline
#endif

namespace dcd {

struct CudaError : std::runtime_error {
    explicit CudaError(const std::string& s) : std::runtime_error(s) {}
};
struct StatusError : std::runtime_error {
    int code;
    StatusError(int c, const std::string& s) : std::runtime_error(s), code(c) {}
};

#define DCD_CUDA(expr)                                                                                   \
    do {                                                                                                 \
        cudaError_t _e = (expr);                                                                         \
        if (_e != cudaSuccess)                                                                           \
            throw dcd::CudaError(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " + __FILE__ + ":" + \
                                 std::to_string(__LINE__));                                              \
    } while (0)

// Packed posting key: codeId in the high 24 bits, coordinate in the low 40.
constexpr int COORD_BITS = 40;
constexpr u64 COORD_MASK = (1ull << COORD_BITS) - 1;
constexpr int MAX_KEY_UNITS = 48;   // longest string held by the device string table
constexpr int MAX_CODES = 6;        // self + up to 4 vocabularies + "$stem"
constexpr int MAX_VOCS = 4;

// character class bits (same meaning as tools/gen_unicode_tables.py)
enum { CLS_L = 1, CLS_N = 2, CLS_W = 4 };

// entry flags of the string table
enum : u32 {
    EF_WORD = 1u,      // seen as a token / AddWord argument: needs GetWordCodes
    EF_CODE = 2u,      // used as an index key (gets an ordinal code id)
    EF_STOP = 4u,      // member of the stop list
    EF_LOCAL = 8u,     // sharded build: code present on THIS rank (the table also holds the other ranks' codes)
};

#if defined(__CUDACC__)
DCD_D u32 lane_id() { return threadIdx.x & 31; }
DCD_D u32 warp_id() { return threadIdx.x >> 5; }

DCD_D u64 ld_volatile_u64(const u64* p) {
    u64 v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
DCD_D u32 ld_volatile_u32(const u32* p) {
    u32 v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
// acquire load / release store at gpu scope: publication of a record through one flag word without a full __threadfence()
// on either side (MEMBAR.SC showed up as the top stall of the string-table kernels)
DCD_D u32 ld_acquire_u32(const u32* p) {
    u32 v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
DCD_D void st_release_u32(u32* p, u32 v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
DCD_D void st_volatile_u64(u64* p, u64 v) { asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
DCD_D void st_volatile_u32(u32* p, u32 v) { asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }

// streaming 128-bit load that does not pollute L1 (text and keys are read exactly once per pass)
DCD_D uint4 ld_stream_u128(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

template <typename T>
DCD_D T warp_incl_scan(T v) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        T t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane_id() >= (u32)o) v += t;
    }
    return v;
}
template <typename T>
DCD_D T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide exclusive scan for blockDim.x == NT (multiple of 32, <= 1024).  `smem` needs NT/32+1 entries.
// Returns the exclusive prefix of v; *total receives the block total.  Contains two __syncthreads().
template <typename T, int NT>
DCD_D T block_excl_scan(T v, T* smem, T* total) {
    T inc = warp_incl_scan(v);
    if (lane_id() == 31) smem[warp_id()] = inc;
    __syncthreads();
    if (warp_id() == 0) {
        T w = lane_id() < NT / 32 ? smem[lane_id()] : T(0);
        T wi = warp_incl_scan(w);
        if (lane_id() < NT / 32) smem[lane_id()] = wi - w;
        if (lane_id() == NT / 32 - 1) smem[NT / 32] = wi;
    }
    __syncthreads();
    T r = inc - v + smem[warp_id()];
    *total = smem[NT / 32];
    return r;
}

// ---- decoupled look-back (single-pass chained scan) ---------------------------------------------
// status word: [63:62] flag (0 = not ready, 1 = tile aggregate, 2 = inclusive prefix), [61:0] value.
constexpr u64 LB_AGG = 1ull << 62;
constexpr u64 LB_PREFIX = 2ull << 62;
constexpr u64 LB_VALUE_MASK = (1ull << 62) - 1;

// Called by ONE full warp (all 32 lanes).  Publishes this tile's aggregate, walks predecessors, publishes the
// inclusive prefix and returns the exclusive prefix.  Tiles must obtain their index from an atomic ticket so
// that every predecessor is already resident (forward progress).
DCD_D u64 lookback_warp_wait(u64* status, int tile, u64 aggregate);
DCD_D u64 lookback_warp(u64* status, int tile, u64 aggregate) {
    const u32 lane = lane_id();
    if (tile == 0) {
        if (lane == 0) st_volatile_u64(status, LB_PREFIX | aggregate);
        return 0;
    }
    if (lane == 0) st_volatile_u64(status + tile, LB_AGG | aggregate);
    return lookback_warp_wait(status, tile, aggregate);
}
// Second half of lookback_warp for callers that publish (LB_AGG | aggregate, or LB_PREFIX for tile 0) themselves earlier
// so that independent work overlaps the latency of the walk.
DCD_D u64 lookback_warp_wait(u64* status, int tile, u64 aggregate) {
    const u32 lane = lane_id();
    if (tile == 0) return 0;
    u64 exclusive = 0;
    int base = tile - 1;
    while (true) {
        int j = base - (int)lane;
        u64 s = j >= 0 ? ld_volatile_u64(status + j) : LB_PREFIX;
        while (__any_sync(0xffffffffu, (s >> 62) == 0)) {
            if ((s >> 62) == 0) s = ld_volatile_u64(status + j);
        }
        u32 pm = __ballot_sync(0xffffffffu, (s >> 62) == 2);
        int first = pm ? __ffs(pm) - 1 : 32;
        u64 v = ((int)lane <= first) ? (s & LB_VALUE_MASK) : 0;
        exclusive += warp_sum(v);
        if (pm) break;
        base -= 32;
    }
    if (lane == 0) st_volatile_u64(status + tile, LB_PREFIX | (exclusive + aggregate));
    return exclusive;
}
#endif  // __CUDACC__

}  // namespace dcd
