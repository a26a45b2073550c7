// Single-pass tokenizer, word-table listing and posting expansion for sm_100a (see tokenize.cuh).
#include "tokenize.cuh"
#include "sort.cuh"

#include <algorithm>

namespace dcd {

// ---------------------------------------------------------------------------------------------------
// per-tile page info: first page of the tile and the common coordinate delta of its pages
// ---------------------------------------------------------------------------------------------------
__global__ void k_tile_info(const u64* __restrict__ page_off, const i64* __restrict__ page_delta, u64 npages, u64 n_units,
                            u32* __restrict__ first_page, i64* __restrict__ tile_delta, u32 ntiles) {
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const u64 x = (u64)t * TK_TILE;
    u64 lo = 0, hi = npages;   // last p with page_off[p] <= x  (page_off[0] == 0)
    while (hi - lo > 1) {
        const u64 mid = (lo + hi) >> 1;
        if (page_off[mid] <= x) lo = mid; else hi = mid;
    }
    // empty pages share their offset with the next page: start from the last page at this offset
    first_page[t] = (u32)lo;
    const u64 end = min(x + (u64)TK_TILE, n_units);
    i64 d = page_delta[lo];
    bool uniform = d != PAGE_SKIP;
    for (u64 p = lo + 1; uniform && p < npages && page_off[p] < end; p++) uniform = page_delta[p] == d;
    tile_delta[t] = uniform ? d : TILE_DELTA_MIXED;
}
void launch_tile_info(const u64* page_off, const i64* page_delta, u64 npages, u64 n_units, u32* tile_first_page, i64* tile_delta,
                      u32 ntiles, cudaStream_t s) {
    if (!ntiles) return;
    k_tile_info<<<(ntiles + 127) / 128, 128, 0, s>>>(page_off, page_delta, npages, n_units, tile_first_page, tile_delta, ntiles);
}

// page-start bitmask: bit x set iff some page starts at unit offset x (bits, not a per-tile search: the tokenizer
// reads its 32-unit word of the mask together with its 32 units of text)
__global__ void k_page_bits(const u64* __restrict__ page_off, u64 npages, u32* __restrict__ bits) {
    const u64 p = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npages) return;
    const u64 off = page_off[p];
    if (p > 0 && page_off[p - 1] == off) return;   // empty predecessor: same bit
    atomicOr(bits + (off >> 5), 1u << (off & 31));
}
void launch_page_bits(const u64* page_off, u64 npages, u32* bits, cudaStream_t s) {
    if (!npages) return;
    k_page_bits<<<(unsigned)((npages + 255) / 256), 256, 0, s>>>(page_off, npages, bits);
}

// ---------------------------------------------------------------------------------------------------
// classification: 4 ASCII bytes at a time
// ---------------------------------------------------------------------------------------------------
// bit 7 of every byte -> 4-bit nibble (byte j -> bit j)
DCD_D u32 msb_nibble(u32 m) { return ((m >> 7) * 0x10204080u) >> 28; }

// w: 4 text bytes, all < 0x80.  Returns the lower-cased bytes; letter / digit nibbles through references.
DCD_D u32 classify_ascii4(u32 w, u32& nl, u32& nd) {
    const u32 up = ((w + 0x3F3F3F3Fu) & ~(w + 0x25252525u)) & 0x80808080u;   // 'A'..'Z'
    const u32 lw = w | (up >> 2);
    nl = msb_nibble(((lw + 0x1F1F1F1Fu) & ~(lw + 0x05050505u)) & 0x80808080u);   // 'a'..'z'
    nd = msb_nibble(((w + 0x50505050u) & ~(w + 0x46464646u)) & 0x80808080u);     // '0'..'9'
    return lw;
}

// one UTF-16 unit (or Latin-1 byte) through the Unicode tables: returns the lower-cased unit, class in k (1 letter, 2 number)
DCD_D u32 classify_unit(u32 c, const u8* __restrict__ cls, const u16* __restrict__ lower, u32& k) {
    u32 l;
    if (c < 128) {
        l = (c >= 'A' && c <= 'Z') ? c + 32 : c;
        k = ((l >= 'a' && l <= 'z') ? 1u : 0u) | ((c >= '0' && c <= '9') ? 2u : 0u);
    } else {
        l = __ldg(lower + c);
        const u32 b = __ldg(cls + l);
        k = (b & CLS_L) ? 1u : ((b & CLS_N) ? 2u : 0u);
    }
    return l;
}

// 4 Latin-1 bytes (any value): lower-cased bytes + class nibbles
DCD_D u32 classify_latin4(u32 w, const u8* __restrict__ cls, const u16* __restrict__ lower, u32& nl, u32& nd) {
    if ((w & 0x80808080u) == 0) return classify_ascii4(w, nl, nd);
    u32 out = 0;
    nl = nd = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        u32 k;
        const u32 l = classify_unit((w >> (8 * j)) & 0xFF, cls, lower, k);
        out |= (l & 0xFF) << (8 * j);   // ToLower of a Latin-1 letter stays in Latin-1
        nl |= (k & 1u) << j;
        nd |= (k >> 1) << j;
    }
    return out;
}

template <int ENC>
DCD_D u32 load_text_unit(const void* text, u64 x) {
    return ENC == 1 ? (u32)__ldg((const u8*)text + x) : (u32)__ldg((const u16*)text + x);
}

// Key byte of a lower-cased UTF-16 unit (wordtable.cuh): the unit itself below U+0100, 0x80 | low 7 bits inside the Cyrillic
// block, 0 for anything else; kind: 1 Latin-1 upper half, 2 Cyrillic block, 4 not representable
DCD_D u32 key_byte(u32 l, u32& kind) {
    if (l < 0x80u) { kind = 0; return l; }
    if (l < 0x100u) { kind = 1; return l; }
    if (is_cyr_block(l)) { kind = 2; return 0x80u | (l & 0x7Fu); }
    kind = 4;
    return 0;
}

// Units the exact keys cannot hold, per 32-unit chunk (UTF-16 input only; Latin-1 input is always tag "bytes")
struct XMask { u32 x1, x3, xb; };   // Latin-1 upper half, Cyrillic block, neither

// 32 units at unit offset x0: lower-cased KEY BYTES to ch8 (the lower-cased units to ch16 for UTF-16 input), class masks out
template <int ENC>
DCD_D void classify32(const void* text, u64 n_units, u64 x0, const u8* __restrict__ cls, const u16* __restrict__ lower, u8* ch8,
                      u16* ch16, u32& L, u32& D, XMask& X) {
    L = D = 0;
    X.x1 = X.x3 = X.xb = 0;
    if (x0 + 32 <= n_units) {
        if (ENC == 1) {
            const uint4* p = reinterpret_cast<const uint4*>((const u8*)text + x0);
            const uint4 a = __ldg(p), b = __ldg(p + 1);
            const u32 w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
            u32 lw[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                u32 nl, nd;
                lw[j] = classify_latin4(w[j], cls, lower, nl, nd);
                L |= nl << (4 * j);
                D |= nd << (4 * j);
            }
            uint4* d = reinterpret_cast<uint4*>(ch8);
            d[0] = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            d[1] = make_uint4(lw[4], lw[5], lw[6], lw[7]);
        } else {
            const uint4* p = reinterpret_cast<const uint4*>((const u16*)text + x0);
            u32 sh[8];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint4 v = __ldg(p + q);
                const u32 ww[4] = {v.x, v.y, v.z, v.w};
                u32 o[4];
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    u32 k0, k1, t0, t1;
                    const u32 l0 = classify_unit(ww[r] & 0xFFFF, cls, lower, k0);
                    const u32 l1 = classify_unit(ww[r] >> 16, cls, lower, k1);
                    o[r] = l0 | (l1 << 16);
                    const int j = q * 8 + r * 2;
                    L |= ((k0 & 1u) << j) | ((k1 & 1u) << (j + 1));
                    D |= ((k0 >> 1) << j) | ((k1 >> 1) << (j + 1));
                    const u32 s0 = key_byte(l0, t0), s1 = key_byte(l1, t1);
                    X.x1 |= ((t0 & 1u) << j) | ((t1 & 1u) << (j + 1));
                    X.x3 |= (((t0 >> 1) & 1u) << j) | (((t1 >> 1) & 1u) << (j + 1));
                    X.xb |= ((t0 >> 2) << j) | ((t1 >> 2) << (j + 1));
                    const u32 pair = s0 | (s1 << 8);
                    if (r & 1) sh[q * 2 + (r >> 1)] |= pair << 16; else sh[q * 2 + (r >> 1)] = pair;
                }
                reinterpret_cast<uint4*>(ch16)[q] = make_uint4(o[0], o[1], o[2], o[3]);
            }
            uint4* d = reinterpret_cast<uint4*>(ch8);
            d[0] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
            d[1] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
        }
    } else {   // tail of the text: scalar, bounds-checked
        for (int j = 0; j < 32; j++) {
            const u64 x = x0 + (u64)j;
            u32 l = 0, k = 0, t = 0;
            if (x < n_units) l = classify_unit(load_text_unit<ENC>(text, x), cls, lower, k);
            if (ENC == 1) ch8[j] = (u8)l;
            else {
                ch16[j] = (u16)l;
                ch8[j] = (u8)key_byte(l, t);
                X.x1 |= (t & 1u) << j;
                X.x3 |= ((t >> 1) & 1u) << j;
                X.xb |= (t >> 2) << j;
            }
            L |= (k & 1u) << j;
            D |= (k >> 1) << j;
        }
    }
}

// Break mask of 32 units: bit j set iff unit j does NOT continue the run of unit j - 1 (other class, no class, or a page
// starts at j).  pl / pd: class of the unit in front of the 32.  Token starts are C & brk; the token that starts at bit i ends
// in front of the first break behind i.
DCD_D u32 break_mask(u32 L, u32 D, u32 P, u32 pl, u32 pd) {
    const u32 cont = ((L & ((L << 1) | pl)) | (D & ((D << 1) | pd))) & ~P;
    return ~cont;
}
// starts of the tokens with at least min_len units among the starts S; brk64 = break bits of these 32 units and the next 32
DCD_D u32 long_enough(u32 S, u64 brk64, int min_len) {
    const u64 cont64 = ~brk64;
    u32 V = S;
    for (int j = 1; j < min_len && j <= 32; j++) V &= (u32)(cont64 >> j);
    return V;
}
// length of the token that starts at bit i (1..32), 0 when it is longer than max_len (the whole match is skipped, Build.cs:528)
DCD_D int token_len(u64 brk64, int i, int max_len) {
    const u64 x = brk64 >> (i + 1);
    const int len = x ? __ffsll((long long)x) : 99;
    return len > max_len ? 0 : len;
}

// Find-or-insert the token at ch8[start, start + len): an exact raw-byte key when it has at most 15 units the key bytes can
// hold (two funnel shifts and a mask), else the generic path over the lower-cased units.  xm: XMask per 32-unit chunk of
// the tile (UTF-16 input; unused for Latin-1 input, whose bytes are always tag "bytes").
// 4 lower-case ASCII letters (or zero padding) -> 20 bits, 5 per letter
DCD_D u32 pack_alpha4(u32 v) {
    const u32 x = v & 0x1F1F1F1Fu;
    const u32 t = (x & 0x001F001Fu) | ((x & 0x1F001F00u) >> 3);
    return (t & 0x3FFu) | ((t >> 6) & 0xFFC00u);
}

template <int ENC>
DCD_D u32 intern_token(const WordTable& tab, const u8* ch8, const u16* ch16, const XMask* xm, int start, int len, bool dig) {
    u32 tag = WK_BYTES;
    bool exact = len <= WK_MAX_EXACT;
    bool ascii = true;
    if (ENC == 2 && len <= WK_MAX_ALPHA) {
        const int c = start >> 5, i = start & 31;
        const u64 lm = (1ull << len) - 1;
        const u64 b = (((u64)xm[c + 1].xb << 32 | xm[c].xb) >> i) & lm;
        const u64 h = (((u64)xm[c + 1].x1 << 32 | xm[c].x1) >> i) & lm;
        const u64 y = (((u64)xm[c + 1].x3 << 32 | xm[c].x3) >> i) & lm;
        ascii = !(b | h | y);
        if (b || (h && y)) exact = false;
        else if (y) tag = WK_CYR;
    }
    if (!exact && !dig && len <= WK_MAX_ALPHA && ascii) {   // 16..24 letters: 5-bit packing when they are all a-z
        const u32* w32 = reinterpret_cast<const u32*>(ch8 + (start & ~3));
        const int shb = (start & 3) * 8;
        u64 lo = 0, hi = 0;
        u32 hb = 0;
#pragma unroll
        for (int j = 0; j < 6; j++) {
            u32 v = __funnelshift_r(w32[j], w32[j + 1], shb);
            const int rem = len - 4 * j;
            if (rem <= 0) v = 0;
            else if (rem < 4) v &= (1u << (8 * rem)) - 1u;
            hb |= v;
            const u64 p = pack_alpha4(v);
            if (j < 3) lo |= p << (20 * j);
            else if (j == 3) { lo |= p << 60; hi |= p >> 4; }
            else hi |= p << (20 * j - 64);
        }
        if (!(hb & 0x80808080u)) return wordtab_intern(tab, lo, hi | ((u64)WK_ALPHA << WK_TAG_SHIFT));
    }
    if (exact) {
        const u32* w32 = reinterpret_cast<const u32*>(ch8 + (start & ~3));
        const int shb = (start & 3) * 8;
        const u32 a0 = w32[0], a1 = w32[1], a2 = w32[2];
        u64 lo = (u64)__funnelshift_r(a0, a1, shb) | ((u64)__funnelshift_r(a1, a2, shb) << 32);
        u64 hi = 0;
        if (len < 8) lo &= (1ull << (8 * len)) - 1;
        else if (len > 8) {
            const u32 a3 = w32[3], a4 = w32[4];
            hi = ((u64)__funnelshift_r(a2, a3, shb) | ((u64)__funnelshift_r(a3, a4, shb) << 32)) & ((1ull << (8 * (len - 8))) - 1);
        }
        return wordtab_intern(tab, lo, hi | ((u64)tag << WK_TAG_SHIFT));
    }
    if (ENC == 1) return wordtab_intern_units(tab, (const u8*)(ch8 + start), len);
    return wordtab_intern_units(tab, (const u16*)(ch16 + start), len);
}

struct alignas(16) TokSmemHdr {
    u32 mL[TK_THREADS + 4];   // [0]: bit 31 = class of the unit before the tile; [1..256] threads; [257,258] look-ahead
    u32 mD[TK_THREADS + 4];
    u32 mB[TK_THREADS + 4];   // break masks, [1 + chunk]
    XMask xm[TK_THREADS + 4]; // [chunk]
    u32 scan[TK_THREADS / 32 + 1];
    u32 ntok;
    u32 pad[3];
};

constexpr int CH8_BYTES = TK_TILE + TK_LA + 64;            // tile + look-ahead + window over-read pad
constexpr int CH16_BYTES = (TK_TILE + TK_LA) * 2;

size_t tokenize_smem_bytes(int enc, u32 tok_cap) {
    return sizeof(TokSmemHdr) + CH8_BYTES + (enc == 2 ? CH16_BYTES : 0) + (size_t)tok_cap * 4;
}

// v1 (explicit pages, per-page coordinate deltas): one CTA per 8192-unit tile, 32 units per thread
template <int ENC>   // 1 = Latin-1 bytes, 2 = UTF-16 units
__global__ void __launch_bounds__(TK_THREADS) k_tokenize(TextView tv, TokenizeParams tp) {
    extern __shared__ __align__(16) unsigned char smraw[];
    TokSmemHdr& sm = *reinterpret_cast<TokSmemHdr*>(smraw);
    u8* ch8 = smraw + sizeof(TokSmemHdr);
    u16* ch16 = reinterpret_cast<u16*>(ch8 + CH8_BYTES);
    u32* tokl = reinterpret_cast<u32*>(ch8 + CH8_BYTES + (ENC == 2 ? CH16_BYTES : 0));
    const int tid = threadIdx.x;
    u64 my_tokens = 0;
    for (int i = TK_TILE + TK_LA + tid; i < CH8_BYTES; i += TK_THREADS) ch8[i] = 0;

    for (u32 tile = blockIdx.x; tile < tv.ntiles; tile += gridDim.x) {
        const u64 t0 = (u64)tile * TK_TILE;
        __syncthreads();   // the previous tile's token phase is done with ch8 / tokl / masks

        // ---- 1. load, lower-case, classify: chunk c covers units [32c, 32c + 32); chunks TK_THREADS, TK_THREADS + 1 = look-ahead
        for (int c = tid; c < TK_THREADS + TK_LA / TK_CPT; c += TK_THREADS) {
            u32 L, D;
            XMask X;
            classify32<ENC>(tv.text, tv.n_units, t0 + (u64)c * TK_CPT, tp.cls, tp.lower, ch8 + c * TK_CPT, ch16 + c * TK_CPT, L, D, X);
            sm.mL[1 + c] = L;
            sm.mD[1 + c] = D;
            sm.xm[c] = X;
        }
        if (tid == 32) {   // the unit before the tile
            u32 k = 0;
            if (t0 > 0) classify_unit(load_text_unit<ENC>(tv.text, t0 - 1), tp.cls, tp.lower, k);
            sm.mL[0] = (k & 1u) << 31;
            sm.mD[0] = (k >> 1) << 31;
        }
        __syncthreads();
        for (int c = tid; c < TK_THREADS + TK_LA / TK_CPT; c += TK_THREADS)
            sm.mB[1 + c] = break_mask(sm.mL[1 + c], sm.mD[1 + c], __ldg(tv.page_bits + (t0 >> 5) + c), sm.mL[c] >> 31, sm.mD[c] >> 31);
        __syncthreads();

        // ---- 2. token starts and lengths by bit arithmetic ---------------------------------------------
        const u32 D = sm.mD[1 + tid];
        const u32 B = sm.mB[1 + tid];
        const u64 brk64 = ((u64)sm.mB[2 + tid] << 32) | B;
        const u32 V = long_enough((sm.mL[1 + tid] | D) & B, brk64, tp.min_len);
        u32 ntok;
        u32 off = block_excl_scan<u32, TK_THREADS>((u32)__popc(V), sm.scan, &ntok);
        for (u32 s = V; s; s &= s - 1) {
            const int i = __ffs(s) - 1;
            const int len = token_len(brk64, i, tp.max_len);
            tokl[off++] = (u32)(tid * TK_CPT + i) | ((u32)len << 13) | (((D >> i) & 1u) << 19);
        }
        __syncthreads();

        // ---- 3. per token: exact key, one probe, token record --------------------------------------------
        const i64 tdelta = tv.tile_delta[tile];
        u64* out = tp.tok + (u64)tile * tp.tok_cap;
        for (u32 k = tid; k < ntok; k += TK_THREADS) {
            const u32 e = tokl[k];
            const int start = (int)(e & 0x1FFF), len = (int)((e >> 13) & 63);
            const u64 x = t0 + (u64)start;
            i64 delta = tdelta;
            if (len == 0) delta = PAGE_SKIP;
            else if (tdelta == TILE_DELTA_MIXED) {   // pages of this tile have different origins (explicit pages / header pages)
                u64 lo = tv.tile_first_page[tile], hi = tv.npages;    // last page with page_off <= x
                while (hi - lo > 1) {
                    const u64 mid = (lo + hi) >> 1;
                    if (tv.page_off[mid] <= x) lo = mid; else hi = mid;
                }
                delta = tv.page_delta[lo];
            }
            u32 slot = WORD_SLOT_NONE;
            u64 coord = 0;
            if (delta != PAGE_SKIP) {
                coord = (u64)((i64)x + delta);
                if (coord > COORD_MASK) atomicExch(tp.error + 2, 1u);
                slot = intern_token<ENC>(tp.tab, ch8, ch16, sm.xm, start, len, (e >> 19) & 1u);
                my_tokens++;
            }
            out[k] = ((u64)slot << COORD_BITS) | (coord & COORD_MASK);
        }
        if (tid == 0) tp.tile_count[tile] = ntok;
    }
    my_tokens = warp_sum(my_tokens);
    if (lane_id() == 0 && my_tokens) atomicAdd(tp.token_total, my_tokens);
}

void launch_tokenize(const TextView& tv, const TokenizeParams& tp, int nsm, cudaStream_t s) {
    if (!tv.ntiles) return;
    const size_t smem = tokenize_smem_bytes(tv.enc, tp.tok_cap);
    const void* fn = tv.enc == 1 ? (const void*)k_tokenize<1> : (const void*)k_tokenize<2>;
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, TK_THREADS, smem);
    if (occ < 1) occ = 1;
    const unsigned grid = (unsigned)std::min<u64>((u64)nsm * occ, tv.ntiles);
    if (tv.enc == 1) k_tokenize<1><<<grid, TK_THREADS, smem, s>>>(tv, tp);
    else k_tokenize<2><<<grid, TK_THREADS, smem, s>>>(tv, tp);
}

// ---------------------------------------------------------------------------------------------------
// occupied slots -> dense work list for K4
// ---------------------------------------------------------------------------------------------------
__global__ void k_list_words(WordTable t, u32* __restrict__ out, u32* __restrict__ out_n) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    bool has = false;
    if (i <= t.cap_mask) {
        const ulonglong2 s = t.slots[i];
        has = (s.x | s.y) != 0 && (s.y >> WK_TAG_SHIFT) != WK_SENTINEL;
    }
    const u32 m = __ballot_sync(0xffffffffu, has);
    u32 base = 0;
    if (lane_id() == 0 && m) base = atomicAdd(out_n, (u32)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (has) out[base + __popc(m & ((1u << lane_id()) - 1))] = i;
}
void launch_list_words(const WordTable& tab, u32* out_slots, u32* out_n, cudaStream_t s) {
    const u64 cap = (u64)tab.cap_mask + 1;
    k_list_words<<<(unsigned)((cap + 255) / 256), 256, 0, s>>>(tab, out_slots, out_n);
}

// ---------------------------------------------------------------------------------------------------
// K3: token records -> posting keys
// ---------------------------------------------------------------------------------------------------
constexpr int EX_THREADS = 256;

DCD_D u32 slot_codes_get(u64 c, const u32* __restrict__ overflow, u32* ids) {   // returns n, fills ids[0..n)
    const u32 n = (u32)(c & 15u);
    if (n <= 2) {
        ids[0] = (u32)(c >> 4) & 0xFFFFFFu;
        ids[1] = (u32)(c >> 28) & 0xFFFFFFu;
    } else {
        const u32* r = overflow + (u64)((c >> 4) & 0xFFFFFFFFull) * 8;
        for (u32 j = 0; j < n && j < (u32)MAX_CODES; j++) ids[j] = r[1 + j];
    }
    return n;
}

// One WARP per token tile: no block barriers on the path.  Pass 1 counts the tile's postings, a warp-level decoupled
// look-back turns that into the tile's output offset, pass 2 re-reads the (L1/L2-resident) records and writes the keys.
__global__ void __launch_bounds__(EX_THREADS) k_expand(ExpandParams ep) {
    extern __shared__ u32 hist[];   // [npasses][WIDE_HIST_STRIDE]
    const int tid = threadIdx.x;
    const u32 lane = lane_id();
    for (int i = tid; i < ep.npasses * WIDE_HIST_STRIDE; i += EX_THREADS) hist[i] = 0;
    __syncthreads();
    while (true) {
        u32 tile = 0;
        if (lane == 0) tile = atomicAdd(ep.ticket, 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= ep.ntiles) break;
        const u32 n = ep.tile_count[tile];
        const u64* tk = ep.tok + (u64)tile * ep.tok_cap;
        u32 cnt = 0;
#pragma unroll 4
        for (u32 k = lane; k < n; k += 32) {
            const u32 slot = (u32)(__ldg(tk + k) >> COORD_BITS);
            if (slot != WORD_SLOT_NONE) cnt += (u32)(__ldg(ep.slot_codes + slot) & 15u);
        }
        const u32 total = warp_sum(cnt);
        const u64 base = lookback_warp(ep.lookback, (int)tile, (u64)total);
        if (tile == ep.ntiles - 1 && lane == 0) ep.total[0] = base + total;
        if (base + total > ep.capacity) {
            if (lane == 0) atomicExch(ep.error + 0, 1u);
            continue;
        }
        u64 run = base;
        for (u32 k0 = 0; k0 < n; k0 += 32) {
            const u32 k = k0 + lane;
            u64 t = 0;
            u32 slot = WORD_SLOT_NONE;
            if (k < n) { t = __ldg(tk + k); slot = (u32)(t >> COORD_BITS); }
            u32 ids[MAX_CODES];
            u32 nc = 0;
            if (slot != WORD_SLOT_NONE) nc = slot_codes_get(__ldg(ep.slot_codes + slot), ep.overflow, ids);
            const u32 incl = warp_incl_scan(nc);
            u64 dst = run + incl - nc;
            const u64 coord = t & COORD_MASK;
            for (u32 c = 0; c < nc; c++) {
                const u32 id = ids[c];
                ep.keys[dst++] = ((u64)id << COORD_BITS) | coord;
                for (int ps = 0; ps < ep.npasses; ps++)
                    atomicAdd(&hist[ps * WIDE_HIST_STRIDE + ((id >> ep.pass_shift[ps]) & ((1u << ep.pass_bits[ps]) - 1))], 1u);
            }
            run += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    __syncthreads();
    for (int i = tid; i < ep.npasses * WIDE_HIST_STRIDE; i += EX_THREADS) {
        const u32 v = hist[i];
        if (v) atomicAdd(ep.hist + i, (u64)v);
    }
}
void launch_expand(const ExpandParams& ep, int nsm, cudaStream_t s) {
    if (!ep.ntiles) return;
    int occ = 1;
    const size_t smem = (size_t)ep.npasses * WIDE_HIST_STRIDE * sizeof(u32);
    cudaFuncSetAttribute(k_expand, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_expand, EX_THREADS, smem);
    if (occ < 1) occ = 1;
    const unsigned grid = (unsigned)std::min<u64>((u64)nsm * occ, ep.ntiles);
    k_expand<<<grid, EX_THREADS, smem, s>>>(ep);
}


// ---------------------------------------------------------------------------------------------------
// v2: warp-autonomous tokenizer (see tokenize.cuh)
// ---------------------------------------------------------------------------------------------------
constexpr int T2_CH8 = T2_WIN + 64;                    // window + over-read pad of the key builder
template <int ENC> struct alignas(16) Tok2WarpSmem {
    u8 ch8[T2_CH8];
    u16 ch16[ENC == 2 ? T2_WIN : 8];
    XMask xm[ENC == 2 ? 34 : 1];
    u32 tokl[T2_CAP + 4];
};

template <int ENC>
__global__ void __launch_bounds__(T2_WARPS * 32, 8) k_tokenize2(Tok2Params tp) {
    extern __shared__ __align__(16) unsigned char smraw[];
    Tok2WarpSmem<ENC>& sm = reinterpret_cast<Tok2WarpSmem<ENC>*>(smraw)[threadIdx.x >> 5];
    const u32 lane = lane_id();
    const u32 gw = blockIdx.x * T2_WARPS + (threadIdx.x >> 5), nw = gridDim.x * T2_WARPS;
    for (int i = T2_WIN + (int)lane; i < T2_CH8; i += 32) sm.ch8[i] = 0;
    if (ENC == 2 && lane < 2) sm.xm[32 + lane] = XMask{0, 0, 0};
    u64 my_tokens = 0;
    for (u32 tile = tp.tile_begin + gw; tile < tp.ntiles; tile += nw) {
        const u64 t0 = (u64)tile * T2_OWN;
        __syncwarp();   // the previous tile's token phase is done with ch8 / tokl
        u32 L, D;
        XMask X;
        classify32<ENC>(tp.text, tp.n_units, t0 + 32u * lane, tp.cls, tp.lower, sm.ch8 + 32 * lane, sm.ch16 + 32 * lane, L, D, X);
        if (ENC == 2) sm.xm[lane] = X;
        const u32 P = __ldg(tp.page_bits + (t0 >> 5) + lane);
        // class of the unit before this lane's 32 (lane 0: the unit before the tile)
        u32 pl = __shfl_up_sync(0xffffffffu, L, 1) >> 31, pd = __shfl_up_sync(0xffffffffu, D, 1) >> 31;
        if (lane == 0) {
            u32 k = 0;
            if (t0 > 0) classify_unit(load_text_unit<ENC>(tp.text, t0 - 1), tp.cls, tp.lower, k);
            pl = k & 1u;
            pd = k >> 1;
        }
        const u32 B = break_mask(L, D, P, pl, pd);
        u32 B1 = __shfl_down_sync(0xffffffffu, B, 1);
        if (lane == 31) B1 = 0xFFFFFFFFu;
        const u64 brk64 = ((u64)B1 << 32) | B;
        const u32 V = lane < T2_OWN / 32 ? long_enough((L | D) & B, brk64, tp.min_len) : 0u;
        const u32 cnt = (u32)__popc(V);
        const u32 incl = warp_incl_scan(cnt);
        const u32 ntok = __shfl_sync(0xffffffffu, incl, 31);
        u32 off = incl - cnt;
        for (u32 sbits = V; sbits; sbits &= sbits - 1) {
            const int i = __ffs(sbits) - 1;
            sm.tokl[off++] = (u32)(lane * 32 + i) | ((u32)token_len(brk64, i, tp.max_len) << 10) | (((D >> i) & 1u) << 16);
        }
        __syncwarp();
        u32* out = tp.rec + (u64)tile * T2_CAP;
        for (u32 k = lane; k < ntok; k += 32) {
            const u32 e = sm.tokl[k];
            const int start = (int)(e & 0x3FF), len = (int)((e >> 10) & 63);
            u32 slot = T2_SLOT_NONE;
            if (len) {
                slot = intern_token<ENC>(tp.tab, sm.ch8, sm.ch16, sm.xm, start, len, (e >> 16) & 1u);
                // WORD_SLOT_NONE: table full (counters[1] is set, the host grows the table and retries).  A real word never
                // sits in slot T2_SLOT_NONE: the engine parks a sentinel in the table's last slot.
                if (slot == WORD_SLOT_NONE) slot = T2_SLOT_NONE;
                my_tokens++;
            }
            out[k] = (slot << T2_OFF_BITS) | (u32)start;
        }
        if (lane == 0) tp.tile_count[tile] = (u16)ntok;
    }
    my_tokens = warp_sum(my_tokens);
    if (lane == 0 && my_tokens) atomicAdd(tp.token_total, my_tokens);
}

void launch_tokenize2(const Tok2Params& tp, int nsm, cudaStream_t s) {
    if (tp.ntiles <= tp.tile_begin) return;
    const size_t smem = (tp.enc == 1 ? sizeof(Tok2WarpSmem<1>) : sizeof(Tok2WarpSmem<2>)) * T2_WARPS;
    const void* fn = tp.enc == 1 ? (const void*)k_tokenize2<1> : (const void*)k_tokenize2<2>;
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, T2_WARPS * 32, smem);
    if (occ < 1) occ = 1;
    const unsigned grid = (unsigned)std::min<u64>((u64)nsm * occ, ((u64)(tp.ntiles - tp.tile_begin) + T2_WARPS - 1) / T2_WARPS);
    if (tp.enc == 1) k_tokenize2<1><<<grid, T2_WARPS * 32, smem, s>>>(tp);
    else k_tokenize2<2><<<grid, T2_WARPS * 32, smem, s>>>(tp);
}


// K3 for v2 records: one warp per group of EX2_GROUP tiles.  Pass 1 counts the group's postings, a warp-level decoupled
// look-back turns that into the output offset, pass 2 re-reads the (L1/L2-resident) records and writes the keys.
__global__ void __launch_bounds__(EX_THREADS, 8) k_expand2(Expand2Params ep) {
    extern __shared__ u32 hist[];   // [npasses][WIDE_HIST_STRIDE]
    const int tid = threadIdx.x;
    const u32 lane = lane_id();
    for (int i = tid; i < ep.npasses * WIDE_HIST_STRIDE; i += EX_THREADS) hist[i] = 0;
    __syncthreads();
    const u32 ngroups = (ep.ntiles + EX2_GROUP - 1) / EX2_GROUP;
    while (true) {
        u32 g = 0;
        if (lane == 0) g = atomicAdd(ep.ticket, 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ngroups) break;
        const u32 tile0 = g * EX2_GROUP, tile1 = min(tile0 + (u32)EX2_GROUP, ep.ntiles);
        u32 cnt = 0;
        for (u32 t = tile0; t < tile1; t++) {
            const u32 n = ep.tile_count[t];
            const u32* tk = ep.rec + (u64)t * T2_CAP;
            for (u32 k = lane; k < n; k += 32) {
                const u32 slot = __ldg(tk + k) >> T2_OFF_BITS;
                if (slot != T2_SLOT_NONE) cnt += (u32)(__ldg(ep.slot_codes + slot) & 15u);
            }
        }
        const u32 total = warp_sum(cnt);
        const u64 base = lookback_warp(ep.lookback, (int)g, (u64)total);
        if (g == ngroups - 1 && lane == 0) ep.total[0] = base + total;
        if (base + total > ep.capacity) {
            if (lane == 0) atomicExch(ep.error + 0, 1u);
            continue;
        }
        u64 run = base;
        for (u32 t = tile0; t < tile1; t++) {
            const u32 n = ep.tile_count[t];
            const u32* tk = ep.rec + (u64)t * T2_CAP;
            const u64 cbase = (u64)((i64)((u64)t * T2_OWN) + ep.delta);
            for (u32 k0 = 0; k0 < n; k0 += 32) {
                const u32 k = k0 + lane;
                u32 r = 0, slot = T2_SLOT_NONE;
                if (k < n) { r = __ldg(tk + k); slot = r >> T2_OFF_BITS; }
                u32 ids[MAX_CODES];
                u32 nc = 0;
                if (slot != T2_SLOT_NONE) nc = slot_codes_get(__ldg(ep.slot_codes + slot), ep.overflow, ids);
                const u32 incl = warp_incl_scan(nc);
                u64 dst = run + incl - nc;
                const u64 coord = (cbase + (r & ((1u << T2_OFF_BITS) - 1))) & COORD_MASK;
                for (u32 c = 0; c < nc; c++) {
                    const u32 id = ids[c];
                    ep.keys[dst++] = ((u64)id << COORD_BITS) | coord;
                    for (int ps = 0; ps < ep.npasses; ps++)
                        atomicAdd(&hist[ps * WIDE_HIST_STRIDE + ((id >> ep.pass_shift[ps]) & ((1u << ep.pass_bits[ps]) - 1))], 1u);
                }
                run += __shfl_sync(0xffffffffu, incl, 31);
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < ep.npasses * WIDE_HIST_STRIDE; i += EX_THREADS) {
        const u32 v = hist[i];
        if (v) atomicAdd(ep.hist + i, (u64)v);
    }
}
void launch_expand2(const Expand2Params& ep, int nsm, cudaStream_t s) {
    if (!ep.ntiles) return;
    int occ = 1;
    const size_t smem = (size_t)ep.npasses * WIDE_HIST_STRIDE * sizeof(u32);
    cudaFuncSetAttribute(k_expand2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_expand2, EX_THREADS, smem);
    if (occ < 1) occ = 1;
    const u64 ngroups = ((u64)ep.ntiles + EX2_GROUP - 1) / EX2_GROUP;
    const unsigned grid = (unsigned)std::min<u64>((u64)nsm * occ, (ngroups + EX_THREADS / 32 - 1) / (EX_THREADS / 32));
    k_expand2<<<grid, EX_THREADS, smem, s>>>(ep);
}

}  // namespace dcd
