// Onesweep LSD radix sort for sm_100a: one read + one write of the keys per digit pass, decoupled look-back per
// digit, warp-level match.any ranking (stable), shared-memory staging so that global stores are coalesced runs.
#include "sort.cuh"

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
