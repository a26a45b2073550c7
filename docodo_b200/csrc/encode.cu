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
};

__global__ void __launch_bounds__(ENC_THREADS) k_encode(const u64* __restrict__ keys, u64 n, u32 ncodes, u16* __restrict__ arena,
                                                         u64 arena_cap, u64* __restrict__ code_off,
                                                         const u64* __restrict__ first_base, u64* lookback, u32* ticket,
                                                         u64* total_out, u32* error) {
    extern __shared__ __align__(16) unsigned char smraw[];
    EncSmem& sm = *reinterpret_cast<EncSmem*>(smraw);
    const int tid = threadIdx.x;
    if (tid == 0) sm.tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const u32 tile = sm.tile;
    const u64 t0 = (u64)tile * ENC_TILE;
    const u64 i0 = t0 + (u64)tid * ENC_ITEMS;

    u64 k[ENC_ITEMS];
    u64 prev = 0;
    bool have_prev = false;
    if (i0 < n) {
        if (i0 + ENC_ITEMS <= n) {
            const uint4* p = reinterpret_cast<const uint4*>(keys + i0);
#pragma unroll
            for (int j = 0; j < ENC_ITEMS / 2; j++) {
                uint4 v = ld_stream_u128(p + j);
                k[2 * j] = (u64)v.x | ((u64)v.y << 32);
                k[2 * j + 1] = (u64)v.z | ((u64)v.w << 32);
            }
        } else {
#pragma unroll
            for (int j = 0; j < ENC_ITEMS; j++) k[j] = i0 + j < n ? keys[i0 + j] : 0;
        }
        if (i0 > 0) { prev = keys[i0 - 1]; have_prev = true; }
    }
    u64 delta[ENC_ITEMS];
    u32 len[ENC_ITEMS];
    u32 mine = 0;
#pragma unroll
    for (int j = 0; j < ENC_ITEMS; j++) {
        len[j] = 0;
        if (i0 + j < n) {
            const u64 code = k[j] >> COORD_BITS, coord = k[j] & COORD_MASK;
            const bool head = !have_prev || (prev >> COORD_BITS) != code;
            u64 last = head ? (first_base ? first_base[code] : 0) : (prev & COORD_MASK);
            delta[j] = coord - last;                    // ulong wrap-around like the reference
            len[j] = (u32)varint_len(delta[j]);
            if (head) len[j] |= 0x100;
            mine += len[j] & 0xFF;
            prev = k[j];
            have_prev = true;
        }
    }
    u32 tile_total;
    u32 off = block_excl_scan<u32, ENC_THREADS>(mine, sm.scan, &tile_total);
    if (warp_id() == 0) {
        u64 ex = lookback_warp(lookback, (int)tile, (u64)tile_total);
        if (lane_id() == 0) sm.tile_base = ex;
    }
    __syncthreads();
    const u64 base = sm.tile_base;
#pragma unroll
    for (int j = 0; j < ENC_ITEMS; j++) {
        if (i0 + j < n) {
            u64 d = delta[j];
            const u32 l = len[j] & 0xFF;
            if (len[j] & 0x100) code_off[k[j] >> COORD_BITS] = base + off;
            for (u32 w = 0; w + 1 < l; w++) { sm.stage[off++] = (u16)(0x8000u | (d & 0x7FFF)); d >>= 15; }
            sm.stage[off++] = (u16)d;
        }
    }
    __syncthreads();
    if (base + tile_total > arena_cap) {
        if (tid == 0) atomicExch(error, 1u);
    } else {
        // coalesced copy: scalar head until 16-byte alignment of the destination, then 128-bit stores
        u16* dst = arena + base;
        const u32 tt = tile_total;
        u32 head = (u32)(((16 - ((uintptr_t)dst & 15)) & 15) / 2);
        if (head > tt) head = tt;
        if ((u32)tid < head) dst[tid] = sm.stage[tid];
        const u32 nvec = (tt - head) / 8;
        for (u32 v = tid; v < nvec; v += ENC_THREADS) {
            const u16* s = sm.stage + head + v * 8;
            uint4 q;
            q.x = (u32)s[0] | ((u32)s[1] << 16);
            q.y = (u32)s[2] | ((u32)s[3] << 16);
            q.z = (u32)s[4] | ((u32)s[5] << 16);
            q.w = (u32)s[6] | ((u32)s[7] << 16);
            *reinterpret_cast<uint4*>(dst + head + v * 8) = q;
        }
        const u32 done = head + nvec * 8;
        if (done + tid < tt) dst[done + tid] = sm.stage[done + tid];
    }
    if (t0 + ENC_TILE >= n && tid == 0) {
        total_out[0] = base + tile_total;
        code_off[ncodes] = base + tile_total;
    }
}

void launch_encode(const u64* keys, u64 n, u32 ncodes, u16* arena, u64 arena_cap, u64* code_off, const u64* first_base,
                   u64* lookback, u32* ticket, u64* total_out, u32* error, cudaStream_t s) {
    if (!n) return;
    cudaFuncSetAttribute(k_encode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncSmem));
    const u64 ntiles = (n + ENC_TILE - 1) / ENC_TILE;
    k_encode<<<(unsigned)ntiles, ENC_THREADS, sizeof(EncSmem), s>>>(keys, n, ncodes, arena, arena_cap, code_off, first_base,
                                                                    lookback, ticket, total_out, error);
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
