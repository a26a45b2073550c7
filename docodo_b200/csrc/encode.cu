#include "encode.cuh"

namespace dcd {

size_t encode_scratch_tiles(u64 n) { return (size_t)((n + ENC_TILE - 1) / ENC_TILE); }

DCD_D int varint_len(u64 d) {   // number of ushorts Builder.Add emits for difference d (IndexSequence.cs:68-81)
    int l = 1;
    while (d > 0x7FFF) { d >>= 15; l++; }
    return l;
}

struct EncSmem {
    u16 stage[ENC_TILE * 5 + 8];
    u32 scan[ENC_THREADS / 32 + 1];
    u32 tile;
    u64 tile_base;
    u64 wmax[ENC_THREADS / 32];
};

// One CTA encodes ENC_TILE consecutive postings.  The tile aggregate is published right after the block scan and the
// look-back (warp 0) overlaps the staging of the ushort stream, which only needs tile-local offsets; the global offset is
// applied at copy-out, which moves aligned 32-bit words (the odd head / tail ushort of a tile is stored on its own, so
// neighbouring tiles never write the same word).
//
// MODE 0: the whole encoder.  MODE 1 (ENC_LENGTHS): lengths only -- code_off, total, max coordinate and, left behind in
// `lookback`, every tile's inclusive ushort prefix; nothing is staged or stored to the arena.  MODE 2 (ENC_PLACED): the
// stream of the tiles [tile_first, tile_first + gridDim.x) only, placed by the prefixes a MODE 1 run left in `lookback` (no
// ticket, no look-back, any tile order, code_off untouched) -- the sharded build encodes owner range by owner range with it
// while the previous range is already travelling to its owner.
template <int MODE>
__global__ void __launch_bounds__(ENC_THREADS, 6) k_encode(const u64* __restrict__ keys, u64 n, u32 ncodes, u16* __restrict__ arena,
                                                         u64 arena_cap, u64* __restrict__ code_off,
                                                         const u64* __restrict__ first_base, const u32* __restrict__ glob,
                                                         const u64* __restrict__ key_base, u64* lookback, u32* ticket,
                                                         u64* total_out, u64* max_coord_out, u32* error, u32 tile_first) {
    extern __shared__ __align__(16) unsigned char smraw[];
    EncSmem& sm = *reinterpret_cast<EncSmem*>(smraw);
    const int tid = threadIdx.x;
    const u32 lane = lane_id();
    if (MODE != ENC_PLACED) {
        if (tid == 0) sm.tile = atomicAdd(ticket, 1u);
        __syncthreads();
    }
    const u32 tile = MODE == ENC_PLACED ? tile_first + blockIdx.x : sm.tile;
    const u64 t0 = (u64)tile * ENC_TILE;
    const u64 i0 = t0 + (u64)tid * ENC_ITEMS;

    u64 k[ENC_ITEMS];
    if (i0 + ENC_ITEMS <= n) {
        const uint4* p = reinterpret_cast<const uint4*>(keys + i0);
#pragma unroll
        for (int j = 0; j < ENC_ITEMS / 2; j++) {
            const uint4 v = __ldg(p + j);
            k[2 * j] = (u64)v.x | ((u64)v.y << 32);
            k[2 * j + 1] = (u64)v.z | ((u64)v.w << 32);
        }
    } else {
#pragma unroll
        for (int j = 0; j < ENC_ITEMS; j++) k[j] = i0 + j < n ? keys[i0 + j] : 0;
    }
    // the posting before this thread's first one: the left neighbour's last key (lane 0: from memory)
    u64 prev = __shfl_up_sync(0xffffffffu, k[ENC_ITEMS - 1], 1);
    if (lane == 0 && i0 > 0 && i0 < n) prev = keys[i0 - 1];
    bool have_prev = i0 > 0;
    const u64 prev0 = prev;
    u64 delta[ENC_ITEMS];
    u32 lens = 0;      // 4 bits per item: varint length 1..5, bit 3 = first posting of its code
    u32 mine = 0;
    u64 cmax = 0;
#pragma unroll
    for (int j = 0; j < ENC_ITEMS; j++) {
        if (i0 + j < n) {
            const u64 code = k[j] >> COORD_BITS, coord = k[j] & COORD_MASK;
            const bool head = !have_prev || (prev >> COORD_BITS) != code;
            // key_base: the first-delta base indexed by the key's own (local) code id -- heads of a tile have consecutive ids, so
            // the loads are sequential instead of two dependent gathers through glob[]
            const u64 last = head ? (key_base ? key_base[code] : (first_base ? first_base[glob ? glob[code] : code] : 0)) : (prev & COORD_MASK);
            delta[j] = coord - last;                    // ulong wrap-around like the reference
            const u32 l = (u32)varint_len(delta[j]);
            lens |= (l | (head ? 8u : 0u)) << (4 * j);
            mine += l;
            cmax = max(cmax, coord);
            prev = k[j];
            have_prev = true;
        }
    }
    u32 tile_total;
    const u32 off0 = block_excl_scan<u32, ENC_THREADS>(mine, sm.scan, &tile_total);
    const bool w0 = warp_id() == 0;
    if (MODE != ENC_PLACED && w0 && lane == 0) st_volatile_u64(lookback + tile, (tile == 0 ? LB_PREFIX : LB_AGG) | (u64)tile_total);
    if (MODE != ENC_LENGTHS) {
        u32 off = off0;
#pragma unroll
        for (int j = 0; j < ENC_ITEMS; j++) {
            const u32 l = (lens >> (4 * j)) & 7u;
            if (l) {
                u64 d = delta[j];
                for (u32 w = 0; w + 1 < l; w++) { sm.stage[off++] = (u16)(0x8000u | (d & 0x7FFF)); d >>= 15; }
                sm.stage[off++] = (u16)d;
            }
        }
    }
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, 16));
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, 8));
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, 4));
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, 2));
    cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, 1));
    if (lane == 0) sm.wmax[warp_id()] = cmax;
    if (MODE == ENC_PLACED) {
        if (tid == 0) sm.tile_base = tile ? (lookback[tile - 1] & LB_VALUE_MASK) : 0;
    } else if (w0) {
        const u64 ex = lookback_warp_wait(lookback, (int)tile, (u64)tile_total);
        if (lane == 0) sm.tile_base = ex;
    }
    __syncthreads();
    const u64 base = sm.tile_base;
    if (MODE != ENC_PLACED) {   // per-code stream starts
        u32 off = off0;
#pragma unroll
        for (int j = 0; j < ENC_ITEMS; j++) {
            const u32 f = (lens >> (4 * j)) & 15u;
            if (f & 8u) {
                // the keys of a sharded build carry dense local code ids; lists are laid out by global ordinal id (glob[])
                const u64 lcode = k[j] >> COORD_BITS;
                const u64 code = glob ? glob[lcode] : lcode;
                code_off[code] = base + off;
                if (first_base && !glob) {   // (with glob the gaps are filled by k_fill_gaps behind the encoder: no dependent loads here)
                    // sharded build: codes of the global dictionary that have no posting on this shard get an empty
                    // list that starts where the next present code starts
                    i64 p = j > 0 ? (i64)(k[j - 1] >> COORD_BITS) : (i0 > 0 ? (i64)(prev0 >> COORD_BITS) : -1);
                    if (glob && p >= 0) p = (i64)glob[p];
                    for (i64 g = p + 1; g < (i64)code; g++) code_off[g] = base + off;
                }
            }
            off += f & 7u;
        }
    }
    if (MODE != ENC_PLACED && tid == 0 && max_coord_out) {
        u64 m = 0;
#pragma unroll
        for (int w = 0; w < ENC_THREADS / 32; w++) m = max(m, sm.wmax[w]);
        if (m > ld_volatile_u64(max_coord_out)) atomicMax(max_coord_out, m);
    }
    const u32 tt = tile_total;
    if (MODE == ENC_LENGTHS) {
    } else if (base + tt > arena_cap) {
        if (tid == 0) atomicExch(error, 1u);
    } else if (tt) {
        u16* dst = arena + base;
        const u32 h = (u32)(base & 1);            // odd start: one ushort on its own, then aligned words
        if (tid == 0 && h) dst[0] = sm.stage[0];
        const u32 nw = (tt - h) >> 1;
        u32* dst32 = reinterpret_cast<u32*>(dst + h);
        const u32* st32 = reinterpret_cast<const u32*>(sm.stage);
        if (h == 0) {
            for (u32 i = tid; i < nw; i += ENC_THREADS) dst32[i] = st32[i];
        } else {
            for (u32 i = tid; i < nw; i += ENC_THREADS) dst32[i] = __funnelshift_r(st32[i], st32[i + 1], 16);
        }
        if (tid == 32 && ((tt - h) & 1u)) dst[tt - 1] = sm.stage[tt - 1];
    }
    if (MODE != ENC_PLACED && t0 + ENC_TILE >= n && tid == 0) {
        total_out[0] = base + tile_total;
        code_off[ncodes] = base + tile_total;
    }
}
// sharded build with dense local ids: the codes of the global dictionary that have no posting on this shard get an empty
// list that starts where the next present code starts.  One thread per PRESENT code fills the absent ones in front of it.
// code_off is preset to ~0 (no list head seen), the encoder's heads have written theirs.
__global__ void k_fill_gaps(u64* code_off, const u32* __restrict__ glob, u32 n_local) {
    const u32 l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n_local) return;
    const u32 g1 = glob[l];
    const u64 v = code_off[g1];
    if (v == ~0ull) return;   // (a dictionary code without a posting here: filled by the next code that has one)
    i64 k = (i64)l - 1;
    while (k >= 0 && code_off[glob[k]] == ~0ull) k--;
    const i64 g0 = k >= 0 ? (i64)glob[k] : -1;
    for (i64 g = g0 + 1; g < (i64)g1; g++) code_off[g] = v;
}
// sharded build: absent codes after the last present one
__global__ void k_fill_tail(const u64* __restrict__ keys, u64 n, u32 ncodes, const u32* __restrict__ glob, u64* __restrict__ code_off) {
    u64 last_code = keys[n - 1] >> COORD_BITS;
    if (glob) last_code = glob[last_code];
    const u64 total = code_off[ncodes];
    for (u64 g = last_code + 1 + (u64)blockIdx.x * blockDim.x + threadIdx.x; g < ncodes; g += (u64)gridDim.x * blockDim.x) code_off[g] = total;
}

void launch_encode(const u64* keys, u64 n, u32 ncodes, u16* arena, u64 arena_cap, u64* code_off, const u64* first_base,
                   const u32* glob, u64* lookback, u32* ticket, u64* total_out, u64* max_coord_out, u32* error, cudaStream_t s,
                   const u64* key_base, u32 n_local) {
    if (!n) return;
    cudaFuncSetAttribute(k_encode<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncSmem));
    if (first_base && glob && n_local) cudaMemsetAsync(code_off, 0xFF, ((size_t)ncodes + 1) * sizeof(u64), s);
    const u64 ntiles = (n + ENC_TILE - 1) / ENC_TILE;
    k_encode<0><<<(unsigned)ntiles, ENC_THREADS, sizeof(EncSmem), s>>>(keys, n, ncodes, arena, arena_cap, code_off, first_base,
                                                                       glob, key_base, lookback, ticket, total_out, max_coord_out, error, 0);
    if (first_base && glob && n_local) k_fill_gaps<<<(n_local + 255) / 256, 256, 0, s>>>(code_off, glob, n_local);
    if (first_base) k_fill_tail<<<64, 256, 0, s>>>(keys, n, ncodes, glob, code_off);
}
// ---- lengths only (sharded build, before the owner split) ----------------------------------------------------------------
// Streaming version of the encoder's first half: no ticket, no look-back, nothing staged.  A tile leaves its ushort total in
// tile_total[tile]; a list head leaves (tile << 32 | offset inside the tile) in code_off[code] (and in the offsets of the
// absent codes in front of it, as k_encode does); k_place_offsets turns those into arena offsets once the tile totals are
// scanned.  The scanned totals are exactly the per-tile prefixes ENC_PLACED reads.
DCD_D u32 varint_len_flat(u64 d) {   // == varint_len, without the loop
    return 1u + (d > 0x7FFFull) + (d > 0x3FFFFFFFull) + (d > 0x1FFFFFFFFFFFull) + (d > 0x0FFFFFFFFFFFFFFFull);
}
__global__ void __launch_bounds__(ENC_THREADS, 5) k_encode_lengths(const u64* __restrict__ keys, u64 n, u64* __restrict__ code_off,
                                                                 const u64* __restrict__ first_base, const u32* __restrict__ glob,
                                                                 const u64* __restrict__ key_base, u64* __restrict__ tile_total,
                                                                 u64* max_coord_out) {
    __shared__ u32 s_scan[ENC_THREADS / 32 + 1];
    __shared__ u64 s_wmax[ENC_THREADS / 32];
    const int tid = threadIdx.x;
    const u32 lane = lane_id();
    const u32 tile = blockIdx.x;
    const u64 i0 = (u64)tile * ENC_TILE + (u64)tid * ENC_ITEMS;
    u64 k[ENC_ITEMS];
    if (i0 + ENC_ITEMS <= n) {
        const uint4* p = reinterpret_cast<const uint4*>(keys + i0);
#pragma unroll
        for (int j = 0; j < ENC_ITEMS / 2; j++) {
            const uint4 v = ld_stream_u128(p + j);
            k[2 * j] = (u64)v.x | ((u64)v.y << 32);
            k[2 * j + 1] = (u64)v.z | ((u64)v.w << 32);
        }
    } else {
#pragma unroll
        for (int j = 0; j < ENC_ITEMS; j++) k[j] = i0 + j < n ? keys[i0 + j] : ~0ull;
    }
    u64 prev = __shfl_up_sync(0xffffffffu, k[ENC_ITEMS - 1], 1);
    if (lane == 0 && i0 > 0 && i0 < n) prev = keys[i0 - 1];
    const u64 prev0 = prev;
    bool have_prev = i0 > 0;
    u32 mine = 0, heads = 0;
    u64 cmax = 0;
    u32 lens = 0;   // 4 bits per item (needed only to place the heads)
#pragma unroll
    for (int j = 0; j < ENC_ITEMS; j++) {
        if (i0 + j < n) {
            const u64 coord = k[j] & COORD_MASK;
            const bool head = !have_prev || ((k[j] ^ prev) >> COORD_BITS) != 0;
            u64 last = prev & COORD_MASK;
            if (head) {
                const u64 code = k[j] >> COORD_BITS;
                last = key_base ? key_base[code] : (first_base ? first_base[glob ? glob[code] : code] : 0);
                heads |= 1u << j;
            }
            const u32 l = varint_len_flat(coord - last);
            lens |= l << (4 * j);
            mine += l;
            cmax = max(cmax, coord);
            prev = k[j];
            have_prev = true;
        }
    }
    u32 total;
    const u32 off0 = block_excl_scan<u32, ENC_THREADS>(mine, s_scan, &total);
    if (tid == 0) tile_total[tile] = total;
    if (heads) {
        u32 off = off0;
#pragma unroll
        for (int j = 0; j < ENC_ITEMS; j++) {
            if ((heads >> j) & 1u) {
                const u64 lcode = k[j] >> COORD_BITS;
                const u64 code = glob ? glob[lcode] : lcode;
                const u64 packed = ((u64)tile << 32) | off;
                code_off[code] = packed;
                if (first_base && !glob) {
                    i64 p = j > 0 ? (i64)(k[j - 1] >> COORD_BITS) : (i0 > 0 ? (i64)(prev0 >> COORD_BITS) : -1);
                    if (glob && p >= 0) p = (i64)glob[p];
                    for (i64 g = p + 1; g < (i64)code; g++) code_off[g] = packed;
                }
            }
            off += (lens >> (4 * j)) & 15u;
        }
    }
    if (max_coord_out) {
        for (int o = 16; o > 0; o >>= 1) cmax = max(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
        if (lane == 0) s_wmax[warp_id()] = cmax;
        __syncthreads();
        if (tid == 0) {
            u64 m = 0;
#pragma unroll
            for (int w = 0; w < ENC_THREADS / 32; w++) m = max(m, s_wmax[w]);
            if (m > ld_volatile_u64(max_coord_out)) atomicMax(max_coord_out, m);
        }
    }
}
// code_off[c] (packed by k_encode_lengths) -> arena offset; tile_incl = inclusive scan of tile_total.  Codes behind the last
// head (absent at the tail, or every code of an empty... ) get the total; code_off[ncodes] = total; total_out[0] = total.
__global__ void k_place_offsets(u64* __restrict__ code_off, u32 ncodes, const u64* __restrict__ tile_incl, u64 ntiles, u32 last_code,
                                const u64* __restrict__ keys, u64 n, const u32* __restrict__ glob, u64* __restrict__ total_out) {
    const u64 total = tile_incl[ntiles - 1];
    u64 lc = keys[n - 1] >> COORD_BITS;
    if (glob) lc = glob[lc];
    (void)last_code;
    for (u64 c = (u64)blockIdx.x * blockDim.x + threadIdx.x; c <= ncodes; c += (u64)gridDim.x * blockDim.x) {
        if (c > lc) { code_off[c] = total; continue; }
        const u64 p = code_off[c];
        const u32 t = (u32)(p >> 32);
        code_off[c] = (t ? tile_incl[t - 1] : 0) + (p & 0xFFFFFFFFull);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) total_out[0] = total;
}

void launch_encode_lengths(const u64* keys, u64 n, u32 ncodes, u64* code_off, const u64* first_base, const u32* glob, u64* lookback,
                           u32* ticket, u64* total_out, u64* max_coord_out, cudaStream_t s, const u64* key_base, u32 n_local) {
    if (!n) return;
    // lookback: [2 * ntiles + ntiles / 2048 + 2] u64, zeroed: the scanned prefixes (what ENC_PLACED reads) in front, the raw
    // tile totals behind them, then the scan's own look-back words
    const u64 ntiles = (n + ENC_TILE - 1) / ENC_TILE;
    if (first_base && glob && n_local) cudaMemsetAsync(code_off, 0xFF, ((size_t)ncodes + 1) * sizeof(u64), s);
    u64* tile_total = lookback + ntiles;
    k_encode_lengths<<<(unsigned)ntiles, ENC_THREADS, 0, s>>>(keys, n, code_off, first_base, glob, key_base, tile_total, max_coord_out);
    if (first_base && glob && n_local) k_fill_gaps<<<(n_local + 255) / 256, 256, 0, s>>>(code_off, glob, n_local);   // (packed values)
    launch_inclusive_scan_u64(tile_total, lookback, ntiles, 0, tile_total + ntiles, ticket, s);
    k_place_offsets<<<296, 256, 0, s>>>(code_off, ncodes, lookback, ntiles, 0, keys, n, glob, total_out);
}
size_t encode_lengths_scratch(u64 n) { const u64 t = (n + ENC_TILE - 1) / ENC_TILE; return (size_t)(2 * t + t / 2048 + 4); }
void launch_encode_placed(const u64* keys, u64 n, u64 tile_first, u64 tile_end, u16* arena, u64 arena_cap, const u64* first_base,
                          const u32* glob, const u64* lookback, u32* error, cudaStream_t s, const u64* key_base) {
    if (!n || tile_end <= tile_first) return;
    cudaFuncSetAttribute(k_encode<ENC_PLACED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncSmem));
    k_encode<ENC_PLACED><<<(unsigned)(tile_end - tile_first), ENC_THREADS, sizeof(EncSmem), s>>>(
        keys, n, 0, arena, arena_cap, nullptr, first_base, glob, key_base, const_cast<u64*>(lookback), nullptr, nullptr, nullptr, error,
        (u32)tile_first);
}

// ---------------------------------------------------------------------------------------------------
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = 256 * SCAN_ITEMS;
__global__ void __launch_bounds__(256) k_scan_u64(const u64* __restrict__ in, u64* __restrict__ out, u64 n, u64 base0,
                                                   u64* lookback, u32* ticket) {
    __shared__ u64 sscan[9];
    __shared__ u32 s_tile;
    __shared__ u64 s_base;
    const int tid = threadIdx.x;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const u32 tile = s_tile;
    const u64 i0 = (u64)tile * SCAN_TILE + (u64)tid * SCAN_ITEMS;
    u64 v[SCAN_ITEMS];
    u64 sum = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        v[j] = i0 + j < n ? in[i0 + j] : 0;
        sum += v[j];
    }
    u64 total;
    u64 ex = block_excl_scan<u64, 256>(sum, sscan, &total);
    if (warp_id() == 0) {
        u64 e = lookback_warp(lookback, (int)tile, total);
        if (lane_id() == 0) s_base = e;
    }
    __syncthreads();
    u64 run = base0 + s_base + ex;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; j++) {
        run += v[j];
        if (i0 + j < n) out[i0 + j] = run;
    }
}
void launch_inclusive_scan_u64(const u64* in, u64* out, u64 n, u64 base, u64* lookback, u32* ticket, cudaStream_t s) {
    if (!n) return;
    k_scan_u64<<<(unsigned)((n + SCAN_TILE - 1) / SCAN_TILE), 256, 0, s>>>(in, out, n, base, lookback, ticket);
}

// ---------------------------------------------------------------------------------------------------
// single-list decode: value i ends at the i-th ushort with bit 15 clear; its value is the running sum of the
// (15-bit groups << 15*k) of all groups up to there.  Two-level: per-thread sequential over 8 ushorts needs the
// carry of an unfinished group across threads, so this simple version lets one thread walk a whole value chain
// start; used only for test-sized inputs (dcd_seq_decode).  The batched query decoder lives in query.cu.
__global__ void k_decode_single(const u16* __restrict__ w, u64 n, u64* __restrict__ out, u64* count_out) {
    if (blockIdx.x || threadIdx.x) return;
    u64 last = 0, cnt = 0;
    int shift = 0;
    for (u64 i = 0; i < n; i++) {
        u16 v = w[i];
        if (shift < 64) last += (u64)(v & 0x7FFF) << shift;
        shift += 15;
        if (!(v & 0x8000)) { out[cnt++] = last; shift = 0; }
    }
    count_out[0] = cnt;
}
void launch_decode_single(const u16* words, u64 n, u64* out, u64*, u32*, u64* count_out, cudaStream_t s) {
    k_decode_single<<<1, 32, 0, s>>>(words, n, out, count_out);
}

}  // namespace dcd
