// Onesweep LSD radix sort for sm_100a: one read + one write of the keys per digit pass, decoupled look-back per
// digit, warp-level match.any ranking (stable), shared-memory staging so that global stores are coalesced runs.
#include "sort.cuh"

#include <algorithm>
#include <cstdlib>

namespace dcd {

SortPlan make_sort_plan(int begin_bit, int end_bit) {
    SortPlan p{};
    int bits = end_bit - begin_bit;
    if (bits <= 0) { p.npasses = 0; return p; }
    int np = (bits + 7) / 8;
    p.npasses = np;
    int at = begin_bit;
    for (int i = 0; i < np; i++) {
        int b = bits / np + (i < bits % np ? 1 : 0);
        p.shift[i] = at;
        p.bits[i] = b;
        at += b;
    }
    return p;
}

size_t sort_scratch_bytes(u64 n) {
    u64 ntiles = (n + SORT_TILE - 1) / SORT_TILE;
    return (size_t)(ntiles * 256 * sizeof(u32) + 256 * sizeof(u64) + 64);
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_histogram(const u64* __restrict__ keys, u64 n, SortPlan plan, u64* __restrict__ hist) {
    __shared__ u32 sh[8][256];
    for (int i = threadIdx.x; i < 8 * 256; i += 256) (&sh[0][0])[i] = 0;
    __syncthreads();
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const u64 k = keys[i];
        for (int p = 0; p < plan.npasses; p++) atomicAdd(&sh[p][(k >> plan.shift[p]) & ((1u << plan.bits[p]) - 1)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.npasses * 256; i += 256) {
        u32 v = (&sh[0][0])[i];
        if (v) atomicAdd(hist + i, (u64)v);
    }
}
void launch_histogram(const u64* keys, u64 n, const SortPlan& plan, u64* hist, cudaStream_t s) {
    if (!n || !plan.npasses) return;
    u64 blocks = (n + 256 * 16 - 1) / (256 * 16);
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_histogram<<<(unsigned)blocks, 256, 0, s>>>(keys, n, plan, hist);
}

// exclusive scan of one pass histogram (256 bins) -> global digit bases
__global__ void k_scan_hist(const u64* __restrict__ hist, u64* __restrict__ base) {
    __shared__ u64 sm[9];
    u64 v = hist[threadIdx.x];
    u64 total;
    u64 ex = block_excl_scan<u64, 256>(v, sm, &total);
    base[threadIdx.x] = ex;
}

// status word per (tile, digit): [31:30] flag (1 aggregate, 2 inclusive prefix), [29:0] count
constexpr u32 ST_AGG = 1u << 30, ST_PREFIX = 2u << 30, ST_MASK = (1u << 30) - 1;

template <bool PAIRS>
__global__ void __launch_bounds__(SORT_THREADS) k_onesweep(const u64* __restrict__ in, u64* __restrict__ out,
                                                            const u32* __restrict__ vin, u32* __restrict__ vout, u64 n,
                                                            int shift, int bits, const u64* __restrict__ digit_base,
                                                            u32* status, u32* ticket) {
    extern __shared__ __align__(16) unsigned char smraw[];
    u64* stage = reinterpret_cast<u64*>(smraw);                                   // [SORT_TILE]
    u32* vstage = reinterpret_cast<u32*>(smraw + SORT_TILE * sizeof(u64));        // [SORT_TILE] (PAIRS)
    u32* whist = reinterpret_cast<u32*>(smraw + SORT_TILE * (sizeof(u64) + (PAIRS ? sizeof(u32) : 0)));   // [8][256]
    u32* tile_start = whist + 8 * 256;                                            // [256] digit start within the tile
    u64* gdelta = reinterpret_cast<u64*>(tile_start + 256);                       // [256] global base - tile_start
    __shared__ u32 s_tile;
    __shared__ u32 s_scan[SORT_THREADS / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u32 mask = (1u << bits) - 1;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < 8 * 256; i += SORT_THREADS) whist[i] = 0;
    __syncthreads();
    const u32 tile = s_tile;
    const u64 t0 = (u64)tile * SORT_TILE;
    const int tile_n = (int)min((u64)SORT_TILE, n - t0);

    // 1. load (warp-striped: item i of lane l of warp w is key w*512 + i*32 + l) and rank inside the warp
    u64 key[SORT_ITEMS];
    u32 val[SORT_ITEMS];
    u16 rank[SORT_ITEMS];
    u32* myhist = whist + warp * 256;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int idx = warp * (32 * SORT_ITEMS) + i * 32 + lane;
        if (idx < tile_n) {
            key[i] = in[t0 + idx];
            if (PAIRS) val[i] = vin[t0 + idx];
        } else {
            key[i] = ~0ull;   // padding sorts last inside the (final) tile and is never written
            if (PAIRS) val[i] = 0;
        }
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const u32 d = (u32)(key[i] >> shift) & mask;
        const u32 peers = __match_any_sync(0xffffffffu, d);
        const u32 before = myhist[d];
        rank[i] = (u16)(before + __popc(peers & ((1u << lane) - 1)));
        __syncwarp();
        if ((peers & ((1u << lane) - 1)) == 0) myhist[d] = before + __popc(peers);
        __syncwarp();
    }
    __syncthreads();

    // 2. thread d owns digit d: per-warp exclusive offsets, tile total, look-back
    {
        const int d = tid;
        u32 run = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            u32 c = whist[w * 256 + d];
            whist[w * 256 + d] = run;
            run += c;
        }
        const u32 count = run;
        u32 total;
        const u32 start = block_excl_scan<u32, SORT_THREADS>(count, s_scan, &total);
        tile_start[d] = start;
        // decoupled look-back for this digit
        u32* st = status + (u64)tile * 256 + d;
        u32 excl = 0;
        if (tile == 0) {
            st_volatile_u32(st, ST_PREFIX | count);
        } else {
            st_volatile_u32(st, ST_AGG | count);
            int j = (int)tile - 1;
            while (true) {
                u32 s = ld_volatile_u32(status + (u64)j * 256 + d);
                if ((s >> 30) == 0) continue;
                excl += s & ST_MASK;
                if ((s >> 30) == 2) break;
                j--;
            }
            st_volatile_u32(st, ST_PREFIX | ((excl + count) & ST_MASK));
        }
        gdelta[d] = digit_base[d] + excl - start;
    }
    __syncthreads();

    // 3. scatter into the staging buffer at the tile-local sorted position
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const u32 d = (u32)(key[i] >> shift) & mask;
        const u32 p = tile_start[d] + myhist[d] + rank[i];
        stage[p] = key[i];
        if (PAIRS) vstage[p] = val[i];
    }
    __syncthreads();

    // 4. coalesced write-out: consecutive staging slots of one digit go to consecutive global slots
    for (int p = tid; p < tile_n; p += SORT_THREADS) {
        const u64 k = stage[p];
        const u32 d = (u32)(k >> shift) & mask;
        const u64 dst = gdelta[d] + p;
        out[dst] = k;
        if (PAIRS) vout[dst] = vstage[p];
    }
}

static size_t onesweep_smem(bool pairs) {
    return SORT_TILE * (sizeof(u64) + (pairs ? sizeof(u32) : 0)) + 8 * 256 * sizeof(u32) + 256 * sizeof(u32) + 256 * sizeof(u64);
}

int radix_sort_run(u64* a, u64* b, u32* va, u32* vb, u64 n, const SortPlan& plan, u64* hist, const u64* host_hist,
                   void* scratch, cudaStream_t s, u64* launches) {
    if (!n || !plan.npasses) return 0;
    const bool pairs = va != nullptr;
    const u64 ntiles = (n + SORT_TILE - 1) / SORT_TILE;
    u32* status = (u32*)scratch;
    u64* base = (u64*)((char*)scratch + ntiles * 256 * sizeof(u32));
    u32* ticket = (u32*)(base + 256);
    const size_t smem = onesweep_smem(pairs);
    if (pairs) cudaFuncSetAttribute(k_onesweep<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    else cudaFuncSetAttribute(k_onesweep<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int where = 0;
    for (int p = 0; p < plan.npasses; p++) {
        if (host_hist) {
            bool uniform = false;
            for (int d = 0; d < 256; d++)
                if (host_hist[p * 256 + d] == n) uniform = true;
            if (uniform) continue;
        }
        cudaMemsetAsync(status, 0, ntiles * 256 * sizeof(u32) + 256 * sizeof(u64) + 64, s);
        k_scan_hist<<<1, 256, 0, s>>>(hist + p * 256, base);
        u64* src = where ? b : a;
        u64* dst = where ? a : b;
        u32* vsrc = where ? vb : va;
        u32* vdst = where ? va : vb;
        if (pairs)
            k_onesweep<true><<<(unsigned)ntiles, SORT_THREADS, smem, s>>>(src, dst, vsrc, vdst, n, plan.shift[p], plan.bits[p], base, status, ticket);
        else
            k_onesweep<false><<<(unsigned)ntiles, SORT_THREADS, smem, s>>>(src, dst, nullptr, nullptr, n, plan.shift[p], plan.bits[p], base, status, ticket);
        if (launches) *launches += 2;
        where ^= 1;
    }
    return where;
}

// ---------------------------------------------------------------------------------------------------
// merge a small sorted array into a large one (both sorted by full key); ties: a first
__global__ void k_merge_small(const u64* __restrict__ a, u64 na, const u64* __restrict__ b, u64 nb, u64* __restrict__ out) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na) {
        const u64 k = a[i];
        u64 lo = 0, hi = nb;   // number of b elements < k
        while (lo < hi) {
            u64 mid = (lo + hi) >> 1;
            if (b[mid] < k) lo = mid + 1; else hi = mid;
        }
        out[i + lo] = k;
    } else if (i < na + nb) {
        const u64 j = i - na;
        const u64 k = b[j];
        u64 lo = 0, hi = na;   // number of a elements <= k
        while (lo < hi) {
            u64 mid = (lo + hi) >> 1;
            if (a[mid] <= k) lo = mid + 1; else hi = mid;
        }
        out[lo + j] = k;
    }
}
void launch_merge_small(const u64* a, u64 na, const u64* b, u64 nb, u64* out, cudaStream_t s) {
    u64 n = na + nb;
    if (!n) return;
    k_merge_small<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(a, na, b, nb, out);
}

}  // namespace dcd

// ---------------------------------------------------------------------------------------------------
// Wide-digit onesweep (posting sort).  512 threads x 16 keys = 8192-key tiles, digits of 8..11 bits.
//   1. warp-striped load, match.any ranking against a per-warp u16 running histogram (stable)
//   2. thread t owns DPT consecutive digits: exclusive offsets over the 16 warps, block scan -> tile-local digit
//      starts, per-digit decoupled look-back -> global digit offsets
//   3. keys to their tile-local sorted slot in shared memory (the staging buffer aliases the histograms)
//   4. coalesced write-out: consecutive slots of one digit are consecutive in global memory
// ---------------------------------------------------------------------------------------------------
namespace dcd {

SortPlan make_wide_plan(int begin_bit, int end_bit) {
    SortPlan p{};
    const int bits = end_bit - begin_bit;
    if (bits <= 0) { p.npasses = 0; return p; }
    const int np = (bits + WIDE_MAX_BITS - 1) / WIDE_MAX_BITS;
    p.npasses = np;
    int at = begin_bit;
    for (int i = 0; i < np; i++) {
        const int b = bits / np + (i < bits % np ? 1 : 0);
        p.shift[i] = at;
        p.bits[i] = b;
        at += b;
    }
    return p;
}

// tiles per look-back segment (k_onesweep_wide): 2^17 tiles = 2^30 keys; DCD_SORT_SEG_TILES lowers it so that tests cross
// segment boundaries at test sizes
static u32 wide_seg_tiles() {
    const char* e = getenv("DCD_SORT_SEG_TILES");   // rounded down to a power of two
    u32 v = e ? (u32)std::max(1, std::min(1 << 17, atoi(e))) : (1u << 17);
    while (v & (v - 1)) v &= v - 1;
    return v;
}
static size_t wide_seg_bytes(u64 n, int bits) {      // seg_base: one u64 per (segment, digit)
    const u64 ntiles = (n + WIDE_TILE - 1) / WIDE_TILE;
    return (size_t)((ntiles / wide_seg_tiles() + 1) * ((u64)1 << std::max(bits, 8)) * sizeof(u64));
}
static size_t wide_pass_scratch(u64 n, int bits) {   // status words + digit bases + ticket + segment bases of one pass
    const u64 ntiles = (n + WIDE_TILE - 1) / WIDE_TILE;
    return (size_t)(ntiles * ((u64)1 << std::max(bits, 8)) * sizeof(u32) + WIDE_HIST_STRIDE * sizeof(u64) + 256) + wide_seg_bytes(n, bits);
}
size_t wide_scratch_bytes(u64 n, int npasses) {
    return (size_t)((u64)std::max(npasses, 1) * wide_pass_scratch(n, WIDE_MAX_BITS));
}

__global__ void __launch_bounds__(256) k_histogram_wide(const u64* __restrict__ keys, u64 n, SortPlan plan, u64* __restrict__ hist) {
    extern __shared__ u32 shw[];   // [npasses][WIDE_HIST_STRIDE]
    for (int i = threadIdx.x; i < plan.npasses * WIDE_HIST_STRIDE; i += 256) shw[i] = 0;
    __syncthreads();
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const u64 k = keys[i];
        for (int p = 0; p < plan.npasses; p++) atomicAdd(&shw[p * WIDE_HIST_STRIDE + ((k >> plan.shift[p]) & ((1u << plan.bits[p]) - 1))], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < plan.npasses * WIDE_HIST_STRIDE; i += 256) {
        const u32 v = shw[i];
        if (v) atomicAdd(hist + i, (u64)v);
    }
}
void launch_histogram_wide(const u64* keys, u64 n, const SortPlan& plan, u64* hist, cudaStream_t s) {
    if (!n || !plan.npasses) return;
    u64 blocks = (n + 256 * 16 - 1) / (256 * 16);
    if (blocks > 148 * 4) blocks = 148 * 4;
    const size_t smem = (size_t)plan.npasses * WIDE_HIST_STRIDE * sizeof(u32);
    cudaFuncSetAttribute(k_histogram_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_histogram_wide<<<(unsigned)blocks, 256, smem, s>>>(keys, n, plan, hist);
}

// exclusive scan of one pass histogram (up to 2048 bins) -> global digit bases
__global__ void __launch_bounds__(1024) k_scan_hist_wide(const u64* __restrict__ hist, u64* __restrict__ base) {
    __shared__ u64 sm[33];
    const int t = threadIdx.x;
    const u64 v0 = hist[2 * t], v1 = hist[2 * t + 1];
    u64 total;
    const u64 ex = block_excl_scan<u64, 1024>(v0 + v1, sm, &total);
    base[2 * t] = ex;
    base[2 * t + 1] = ex + v0;
}

// lanes holding the same BITS-bit label, by one ballot per bit (MATCH.ANY runs on the slow ADU pipe: ~57 cycles per
// warp instruction per SM measured on B200, which made it the limiter of the ranking loop)
template <int BITS>
DCD_D u32 match_any_bits(u32 label) {
    u32 peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < BITS; b++) {
        const bool bit = (label >> b) & 1u;
        const u32 m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

template <int N> struct PackU16;
template <> struct PackU16<1> { typedef u16 type; };
template <> struct PackU16<2> { typedef u32 type; };
template <> struct PackU16<4> { typedef u64 type; };

// N consecutive status words read with one volatile load (every word carries its own flag, so a torn vector is harmless)
template <int N> struct StatusVec;
template <> struct StatusVec<1> {
    u32 w[1];
    DCD_D void load(const u32* p) { w[0] = ld_volatile_u32(p); }
};
template <> struct StatusVec<2> {
    u32 w[2];
    DCD_D void load(const u32* p) { asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(w[0]), "=r"(w[1]) : "l"(p)); }
};
template <> struct StatusVec<4> {
    u32 w[4];
    DCD_D void load(const u32* p) {
        asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(p));
    }
};

// Status word of the wide kernel: 0 = not published; bit 31 clear: tile aggregate + 1; bit 31 set: inclusive prefix (31 bits).
constexpr u32 WST_PREFIX = 1u << 31;
// Look-back status words carry 31-bit counts, so the chain is cut into SEGMENTS of 2^17 tiles (2^30 keys): a status
// prefix is relative to the segment's base, seg_base[segment][digit] (u64) = keys of the digit in front of the segment.  The
// first tile of a segment resolves its absolute prefix through the previous segment, publishes the base, then its PREFIX;
// tiles never look back past a segment's first tile (they wait for its PREFIX), so a build is not limited to 2^30 postings.
template <int BITS, bool SEG>
__global__ void __launch_bounds__(WIDE_THREADS, 2) k_onesweep_wide(const u64* __restrict__ in, u64* __restrict__ out, u64 n, int shift,
                                                                    u32 mask, const u64* __restrict__ digit_base, u32* status,
                                                                    u32* ticket, u64* seg_base, u32 seg_shift) {
    constexpr int BINS = 1 << BITS;
    constexpr int NWARP = WIDE_THREADS / 32;
    constexpr int DPT = BINS >= WIDE_THREADS ? BINS / WIDE_THREADS : 1;
    extern __shared__ __align__(16) unsigned char smraw[];
    u64* stage = reinterpret_cast<u64*>(smraw);                     // [WIDE_TILE]   (aliases whist)
    u16* whist = reinterpret_cast<u16*>(smraw);                     // [NWARP][BINS]
    u64* gdelta = reinterpret_cast<u64*>(smraw + WIDE_TILE * sizeof(u64));   // [BINS]
    __shared__ u32 s_tile;
    __shared__ u32 s_scan[NWARP + 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    {   // zero the per-warp histograms (NWARP * BINS u16)
        uint4* z = reinterpret_cast<uint4*>(smraw);
        for (int i = tid; i < NWARP * BINS * 2 / 16; i += WIDE_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    const u32 tile = s_tile;
    const u64 t0 = (u64)tile * WIDE_TILE;
    const int tile_n = (int)min((u64)WIDE_TILE, n - t0);

    // 1. load + rank
    u64 key[WIDE_ITEMS];
    u32 rank2[WIDE_ITEMS / 2];   // two u16 ranks per register
    u16* myhist = whist + warp * BINS;
    const u64* src = in + t0 + warp * (32 * WIDE_ITEMS) + lane;
    if (tile_n == WIDE_TILE) {
#pragma unroll
        for (int i = 0; i < WIDE_ITEMS; i++) key[i] = __ldcs(src + i * 32);
    } else {
#pragma unroll
        for (int i = 0; i < WIDE_ITEMS; i++) {
            const int idx = warp * (32 * WIDE_ITEMS) + i * 32 + lane;
            key[i] = idx < tile_n ? src[i * 32] : ~0ull;   // padding sorts last inside the final tile and is never written
        }
    }
    const u32 lt = (1u << lane) - 1;
#pragma unroll
    for (int i = 0; i < WIDE_ITEMS; i++) {
        const u32 d = (u32)(key[i] >> shift) & mask;
        // ballots keep the ALU pipe busy, MATCH.ANY the ADU pipe: every fourth item goes to the ADU so both pipes work
        const u32 peers = (i % 3) == 2 ? __match_any_sync(0xffffffffu, d) : match_any_bits<BITS>(d);
        const u32 before = myhist[d];
        const u32 r = before + __popc(peers & lt);
        if (i & 1) rank2[i >> 1] |= r << 16; else rank2[i >> 1] = r;
        __syncwarp();
        if ((peers & lt) == 0) myhist[d] = (u16)(before + __popc(peers));
        __syncwarp();
    }
    __syncthreads();

    // 2. digit offsets: thread t owns the DPT consecutive digits [t * DPT, t * DPT + DPT)
    typedef typename PackU16<DPT>::type pack_t;
    u32 cnt[DPT], start[DPT];
    const int d0 = tid * DPT;
    const bool owner = d0 < BINS;
    u32 mine = 0;
#pragma unroll
    for (int j = 0; j < DPT; j++) cnt[j] = 0;
    if (owner) {
#pragma unroll 4
        for (int w = 0; w < NWARP; w++) {
            pack_t* row = reinterpret_cast<pack_t*>(whist + w * BINS + d0);
            const pack_t v = *row;
            pack_t o = 0;
#pragma unroll
            for (int j = 0; j < DPT; j++) {
                o |= (pack_t)cnt[j] << (16 * j);
                cnt[j] += (u32)(v >> (16 * j)) & 0xFFFFu;
            }
            *row = o;
        }
#pragma unroll
        for (int j = 0; j < DPT; j++) mine += cnt[j];
    }
    u32 total;
    u32 run = block_excl_scan<u32, WIDE_THREADS>(mine, s_scan, &total);
    u32* st = status + (u64)tile * BINS + d0;
    if (owner) {
        pack_t sp = 0;
#pragma unroll
        for (int j = 0; j < DPT; j++) { start[j] = run; run += cnt[j]; sp |= (pack_t)start[j] << (16 * j); }
        // publish the tile aggregates now; the look-back itself runs after staging so that its latency overlaps work
#pragma unroll
        for (int j = 0; j < DPT; j++) st_volatile_u32(st + j, tile == 0 ? (WST_PREFIX | cnt[j]) : cnt[j] + 1u);
        // fold the tile-local digit start into the per-warp offsets (sums stay below 2^16: no carry between halves)
#pragma unroll 4
        for (int w = 0; w < NWARP; w++) {
            pack_t* row = reinterpret_cast<pack_t*>(whist + w * BINS + d0);
            *row = *row + sp;
        }
    }
    __syncthreads();

    // 3. tile-local sorted position of every key, then stage (aliases whist: positions first, barrier, then keys)
#pragma unroll
    for (int i = 0; i < WIDE_ITEMS; i++) {
        const u32 d = (u32)(key[i] >> shift) & mask;
        rank2[i >> 1] += (u32)myhist[d] << ((i & 1) * 16);   // positions stay below 2^16: no carry into the upper half
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < WIDE_ITEMS; i++) stage[(rank2[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu] = key[i];

    // 4. decoupled look-back, DPT digits in lock step, LB_K predecessors in flight
    if (owner) {
        u32 excl[DPT];
        u64 base[DPT];
#pragma unroll
        for (int j = 0; j < DPT; j++) { excl[j] = 0; base[j] = 0; }
        if (tile > 0) {
            constexpr int LB_K = 4;
            const u32 seg_mask = (1u << seg_shift) - 1;
            u32 open = (1u << DPT) - 1;   // digits still looking for an inclusive prefix
            int t = (int)tile - 1;
            while (open) {
                StatusVec<DPT> v[LB_K];
#pragma unroll
                for (int q = 0; q < LB_K; q++)
                    if (t - q >= 0) v[q].load(status + (u64)(t - q) * BINS + d0);
#pragma unroll
                for (int q = 0; q < LB_K; q++) {
                    if (!open || t - q < 0) break;
                    const bool seg_first = SEG && ((u32)(t - q) & seg_mask) == 0;   // its aggregate must not be passed: wait for the PREFIX
                    while (true) {
                        bool ready = true;
#pragma unroll
                        for (int j = 0; j < DPT; j++)
                            ready = ready && (!((open >> j) & 1u) || (seg_first ? (v[q].w[j] >> 31) != 0 : v[q].w[j] != 0));
                        if (ready) break;
                        __nanosleep(40);
                        v[q].load(status + (u64)(t - q) * BINS + d0);
                    }
#pragma unroll
                    for (int j = 0; j < DPT; j++) {
                        if ((open >> j) & 1u) {
                            const u32 pf = v[q].w[j] >> 31;          // 1: inclusive prefix, 0: aggregate + 1
                            excl[j] += (v[q].w[j] & ~WST_PREFIX) - 1u + pf;
                            open &= ~(pf << j);
                        }
                    }
                }
                t -= LB_K;
            }
            // every predecessor that was summed lies in the segment of tile - 1: the sums are relative to that segment's base
            const u32 pseg = (tile - 1) >> seg_shift;
            if (SEG && pseg > 0) {
                __threadfence();
#pragma unroll
                for (int j = 0; j < DPT; j++) base[j] = ld_volatile_u64(seg_base + (u64)pseg * BINS + d0 + j);
            }
            if (SEG && (tile & seg_mask) == 0) {   // first tile of a segment: publish the segment's base, then a PREFIX relative to it
#pragma unroll
                for (int j = 0; j < DPT; j++) st_volatile_u64(seg_base + (u64)(tile >> seg_shift) * BINS + d0 + j, base[j] + excl[j]);
                __threadfence();
#pragma unroll
                for (int j = 0; j < DPT; j++) st_volatile_u32(st + j, WST_PREFIX | cnt[j]);
            } else {
#pragma unroll
                for (int j = 0; j < DPT; j++) st_volatile_u32(st + j, WST_PREFIX | (excl[j] + cnt[j]));
            }
        }
#pragma unroll
        for (int j = 0; j < DPT; j++) gdelta[d0 + j] = digit_base[d0 + j] + base[j] + excl[j] - start[j];
    }
    __syncthreads();

    // 5. coalesced write-out
    for (int p = tid; p < tile_n; p += WIDE_THREADS) {
        const u64 k = stage[p];
        const u32 d = (u32)(k >> shift) & mask;
        out[gdelta[d] + p] = k;
    }
}

template <int BITS>
static void launch_wide_pass(const u64* src, u64* dst, u64 n, int shift, int bits, const u64* base, u32* status, u32* ticket,
                             u64* seg_base, cudaStream_t s) {
    const u64 ntiles = (n + WIDE_TILE - 1) / WIDE_TILE;
    const size_t smem = WIDE_TILE * sizeof(u64) + (size_t)(1 << BITS) * sizeof(u64);
    const u32 seg_tiles = wide_seg_tiles();
    if (ntiles > seg_tiles) {   // more than one look-back segment (beyond 2^30 keys): the variant with 64-bit segment bases
        cudaFuncSetAttribute(k_onesweep_wide<BITS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_onesweep_wide<BITS, true><<<(unsigned)ntiles, WIDE_THREADS, smem, s>>>(src, dst, n, shift, (1u << bits) - 1, base, status, ticket,
                                                                               seg_base, (u32)__builtin_ctz(seg_tiles));
    } else {
        cudaFuncSetAttribute(k_onesweep_wide<BITS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_onesweep_wide<BITS, false><<<(unsigned)ntiles, WIDE_THREADS, smem, s>>>(src, dst, n, shift, (1u << bits) - 1, base, status, ticket,
                                                                                seg_base, 31u);
    }
}

int radix_sort_wide(u64* a, u64* b, u64 n, const SortPlan& plan, const u64* hist, void* scratch, cudaStream_t s, u64* launches) {
    if (!n || !plan.npasses) return 0;
    const u64 ntiles = (n + WIDE_TILE - 1) / WIDE_TILE;
    size_t total = 0;
    for (int p = 0; p < plan.npasses; p++) total += wide_pass_scratch(n, plan.bits[p]);
    cudaMemsetAsync(scratch, 0, total, s);
    int where = 0;
    char* sc_next = (char*)scratch;
    for (int p = 0; p < plan.npasses; p++) {
        char* sc = sc_next;
        sc_next += wide_pass_scratch(n, plan.bits[p]);
        u32* status = (u32*)sc;
        u64* base = (u64*)(sc + ntiles * ((u64)1 << std::max(plan.bits[p], 8)) * sizeof(u32));
        u32* ticket = (u32*)(base + WIDE_HIST_STRIDE);
        u64* seg_base = (u64*)(sc + wide_pass_scratch(n, plan.bits[p]) - wide_seg_bytes(n, plan.bits[p]));
        k_scan_hist_wide<<<1, 1024, 0, s>>>(hist + (size_t)p * WIDE_HIST_STRIDE, base);
        const u64* src = where ? b : a;
        u64* dst = where ? a : b;
        const int bits = plan.bits[p];
        if (bits <= 8) launch_wide_pass<8>(src, dst, n, plan.shift[p], bits, base, status, ticket, seg_base, s);
        else if (bits == 9) launch_wide_pass<9>(src, dst, n, plan.shift[p], bits, base, status, ticket, seg_base, s);
        else if (bits == 10) launch_wide_pass<10>(src, dst, n, plan.shift[p], bits, base, status, ticket, seg_base, s);
        else launch_wide_pass<11>(src, dst, n, plan.shift[p], bits, base, status, ticket, seg_base, s);
        if (launches) *launches += 2;
        where ^= 1;
    }
    return where;
}

}  // namespace dcd
