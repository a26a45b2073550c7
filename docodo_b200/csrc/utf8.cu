// UTF-8 -> unit axis of the reference (see utf8.cuh).  sm_100a, warp per 1024-byte block, no block barriers.
#include "utf8.cuh"

#include <algorithm>

namespace dcd {

constexpr int U8_WARPS = 8;

// per 4-byte word: bit 8j set iff byte j is a continuation byte 10xxxxxx
DCD_D u32 cont_bits(u32 w) { return (w >> 7) & ~(w >> 6) & 0x01010101u; }
// bit 8j+4 set iff byte j >= 0xF0 (lead of a 4-byte sequence: two UTF-16 units)
DCD_D u32 four_bits(u32 w) { return w & (w >> 1) & (w >> 2) & (w >> 3) & 0x10101010u; }
// UTF-16 units the bytes of w decode to (w holds 4 valid text bytes)
DCD_D u32 units_of_word(u32 w) { return 4u - (u32)__popc(cont_bits(w)) + (u32)__popc(four_bits(w)); }

// 32 bytes of this lane (zero beyond the end of the text: NUL is one unit, subtracted by the caller through `valid`)
DCD_D void load_lane32(const u8* __restrict__ text, u64 nbytes, u64 x0, u32* w, int& valid) {
    if (x0 + 32 <= nbytes) {
        const uint4 a = ld_stream_u128(text + x0), b = ld_stream_u128(text + x0 + 16);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
        valid = 32;
        return;
    }
    valid = x0 < nbytes ? (int)(nbytes - x0) : 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        u32 v = 0;
        for (int k = 0; k < 4; k++)
            if (4 * j + k < valid) v |= (u32)__ldg(text + x0 + 4 * j + k) << (8 * k);
        w[j] = v;
    }
}

__global__ void __launch_bounds__(U8_WARPS * 32) k_utf8_scan(const u8* __restrict__ text, u64 nbytes, u64 nblocks,
                                                             u64* __restrict__ block_units, u32* __restrict__ flags) {
    const u32 lane = lane_id();
    u32 fl = 0;
    for (u64 blk = (u64)blockIdx.x * U8_WARPS + warp_id(); blk < nblocks; blk += (u64)gridDim.x * U8_WARPS) {
        u32 w[8];
        int valid;
        load_lane32(text, nbytes, blk * U8_BLOCK + 32u * lane, w, valid);
        u32 units = 0, hi = 0, big = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            units += units_of_word(w[j]);
            hi |= w[j] & 0x80808080u;
            big |= ((w[j] & 0x7F7F7F7Fu) + 0x3C3C3C3Cu) & w[j] & 0x80808080u;   // a byte >= 0xC4: code point >= U+0100
        }
        units -= (u32)(32 - valid);   // the zero padding counted one unit per byte
        units = warp_sum(units);
        if (lane == 0) block_units[blk] = units;
        fl |= (hi ? 1u : 0u) | (big ? 2u : 0u);
    }
    fl |= __shfl_xor_sync(0xffffffffu, fl, 16);
    fl |= __shfl_xor_sync(0xffffffffu, fl, 8);
    fl |= __shfl_xor_sync(0xffffffffu, fl, 4);
    fl |= __shfl_xor_sync(0xffffffffu, fl, 2);
    fl |= __shfl_xor_sync(0xffffffffu, fl, 1);
    if (lane == 0 && fl) atomicOr(flags, fl);
}

template <int WIDE>
__global__ void __launch_bounds__(U8_WARPS * 32) k_utf8_transcode(const u8* __restrict__ text, u64 nbytes, u64 nblocks,
                                                                  const u64* __restrict__ block_base, void* __restrict__ out) {
    __shared__ __align__(16) u8 stage[U8_WARPS][U8_BLOCK + 16];
    const u32 lane = lane_id();
    u8* sb = stage[warp_id()];
    for (u64 blk = (u64)blockIdx.x * U8_WARPS + warp_id(); blk < nblocks; blk += (u64)gridDim.x * U8_WARPS) {
        const u64 b0 = blk * U8_BLOCK;
        u32 w[8];
        int valid;
        __syncwarp();
        load_lane32(text, nbytes, b0 + 32u * lane, w, valid);
        uint4* d = reinterpret_cast<uint4*>(sb + 32 * lane);
        d[0] = make_uint4(w[0], w[1], w[2], w[3]);
        d[1] = make_uint4(w[4], w[5], w[6], w[7]);
        if (lane < 3) {   // the first three bytes of the next block (tail of a sequence that starts in this one)
            const u64 x = b0 + U8_BLOCK + lane;
            sb[U8_BLOCK + lane] = x < nbytes ? __ldg(text + x) : (u8)0;
        }
        u32 units = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) units += units_of_word(w[j]);
        units -= (u32)(32 - valid);
        const u32 incl = warp_incl_scan(units);
        u64 o = block_base[blk] + (incl - units);
        __syncwarp();
        const u8* p = sb + 32 * lane;
        for (int j = 0; j < valid; j++) {
            const u32 c = p[j];
            if ((c & 0xC0u) == 0x80u) continue;   // continuation byte: part of the sequence its lead decodes
            u32 cp = c;
            if (c >= 0xF0u) cp = ((c & 0x07u) << 18) | ((p[j + 1] & 0x3Fu) << 12) | ((p[j + 2] & 0x3Fu) << 6) | (p[j + 3] & 0x3Fu);
            else if (c >= 0xE0u) cp = ((c & 0x0Fu) << 12) | ((p[j + 1] & 0x3Fu) << 6) | (p[j + 2] & 0x3Fu);
            else if (c >= 0xC0u) cp = ((c & 0x1Fu) << 6) | (p[j + 1] & 0x3Fu);
            if (WIDE) {
                u16* q = reinterpret_cast<u16*>(out);
                if (c >= 0xF0u) {   // surrogate pair (Encoding.UTF8 -> string)
                    const u32 v = (cp - 0x10000u) & 0xFFFFFu;
                    q[o++] = (u16)(0xD800u + (v >> 10));
                    q[o++] = (u16)(0xDC00u + (v & 0x3FFu));
                } else {
                    q[o++] = (u16)cp;
                }
            } else {
                reinterpret_cast<u8*>(out)[o++] = (u8)cp;
            }
        }
    }
}

__global__ void k_utf8_page_units(const u8* __restrict__ text, u64 nbytes, const u64* __restrict__ block_base, u64 total_units,
                                  const u64* __restrict__ page_off, u64 n, u64* __restrict__ unit_off) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 off = page_off[i];
    if (off >= nbytes) { unit_off[i] = total_units; return; }
    const u64 blk = off / U8_BLOCK, b0 = blk * U8_BLOCK;
    u64 units = block_base[blk];
    u64 x = b0;
    for (; x + 16 <= off; x += 16) {   // b0 is 1024-aligned and text 16-byte aligned
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(text + x));
        units += units_of_word(v.x) + units_of_word(v.y) + units_of_word(v.z) + units_of_word(v.w);
    }
    for (; x < off; x++) {
        const u32 c = __ldg(text + x);
        units += ((c & 0xC0u) != 0x80u ? 1u : 0u) + (c >= 0xF0u ? 1u : 0u);
    }
    unit_off[i] = units;
}

static unsigned warp_grid(u64 nblocks, int nsm) {
    return (unsigned)std::max<u64>(1, std::min<u64>((u64)nsm * 8, (nblocks + U8_WARPS - 1) / U8_WARPS));
}
void launch_utf8_scan(const u8* text, u64 nbytes, u64* block_units, u32* flags, int nsm, cudaStream_t s) {
    const u64 nblocks = (nbytes + U8_BLOCK - 1) / U8_BLOCK;
    if (!nblocks) return;
    k_utf8_scan<<<warp_grid(nblocks, nsm), U8_WARPS * 32, 0, s>>>(text, nbytes, nblocks, block_units, flags);
}
void launch_utf8_transcode(const u8* text, u64 nbytes, const u64* block_base, void* out, int wide, int nsm, cudaStream_t s) {
    const u64 nblocks = (nbytes + U8_BLOCK - 1) / U8_BLOCK;
    if (!nblocks) return;
    if (wide) k_utf8_transcode<1><<<warp_grid(nblocks, nsm), U8_WARPS * 32, 0, s>>>(text, nbytes, nblocks, block_base, out);
    else k_utf8_transcode<0><<<warp_grid(nblocks, nsm), U8_WARPS * 32, 0, s>>>(text, nbytes, nblocks, block_base, out);
}
void launch_utf8_page_units(const u8* text, u64 nbytes, const u64* block_base, u64 total_units, const u64* page_off, u64 n,
                            u64* unit_off, cudaStream_t s) {
    if (!n) return;
    k_utf8_page_units<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(text, nbytes, block_base, total_units, page_off, n, unit_off);
}

}  // namespace dcd
