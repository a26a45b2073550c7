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

struct alignas(16) TokSmemHdr {
    u32 mL[TK_THREADS + 4];   // [0]: bit 31 = class of the unit before the tile; [1..256] threads; [257,258] look-ahead
    u32 mD[TK_THREADS + 4];
    u32 mP[TK_THREADS + 4];   // page starts
    u32 scan[TK_THREADS / 32 + 1];
    u32 ntok;
    u32 pad[3];
};

constexpr int CH8_BYTES = TK_TILE + TK_LA + 64;            // tile + look-ahead + window over-read pad
constexpr int CH16_BYTES = (TK_TILE + TK_LA) * 2;

size_t tokenize_smem_bytes(int enc, u32 tok_cap) {
    return sizeof(TokSmemHdr) + CH8_BYTES + (enc == 2 ? CH16_BYTES : 0) + (size_t)tok_cap * 4;
}

// 4 lower-case ASCII letters (or zero padding) -> 20 bits, 5 per letter
DCD_D u32 pack_alpha4(u32 v) {
    const u32 x = v & 0x1F1F1F1Fu;
    const u32 t = (x & 0x001F001Fu) | ((x & 0x1F001F00u) >> 3);
    return (t & 0x3FFu) | ((t >> 6) & 0xFFC00u);
}
// 4 ASCII digits (or zero padding) -> 16 bits, 4 per digit ('0' -> 1)
DCD_D u32 pack_digit4(u32 v) {
    const u32 x = (v & 0x0F0F0F0Fu) + ((v >> 4) & 0x01010101u);
    const u32 t = (x | (x >> 4)) & 0x00FF00FFu;
    return (t | (t >> 8)) & 0xFFFFu;
}

template <int ENC>
DCD_D u32 load_text_unit(const void* text, u64 x) {
    return ENC == 1 ? (u32)__ldg((const u8*)text + x) : (u32)__ldg((const u16*)text + x);
}

// scalar, bounds-checked classification of 32 units starting at local index `base` (tile tail and look-ahead)
template <int ENC>
DCD_D void classify_span_scalar(const TextView& tv, const TokenizeParams& tp, u64 t0, int base, u8* ch8, u16* ch16, u32& L, u32& D) {
    L = D = 0;
    for (int j = 0; j < 32; j++) {
        const u64 x = t0 + (u64)(base + j);
        u32 l = 0, k = 0;
        if (x < tv.n_units) l = classify_unit(load_text_unit<ENC>(tv.text, x), tp.cls, tp.lower, k);
        if (ENC == 1) ch8[base + j] = (u8)l;
        else { ch16[base + j] = (u16)l; ch8[base + j] = l < 128 ? (u8)l : (u8)0xFF; }
        L |= (k & 1u) << j;
        D |= (k >> 1) << j;
    }
}

// chunk c of the tile starting at unit t0: 32 units -> lower-cased bytes in ch8 (and units in ch16), class masks in sm
template <int ENC>
DCD_D void classify_chunk(const TextView& tv, const TokenizeParams& tp, u64 t0, int c, u8* ch8, u16* ch16, TokSmemHdr& sm) {
    const int base = c * TK_CPT;
    const u64 x0 = t0 + (u64)base;
    u32 L = 0, D = 0;
    if (x0 + TK_CPT <= tv.n_units) {
        if (ENC == 1) {
            const uint4* p = reinterpret_cast<const uint4*>((const u8*)tv.text + x0);
            const uint4 a = __ldg(p), b = __ldg(p + 1);
            const u32 w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
            u32 lw[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                u32 nl, nd;
                lw[j] = classify_latin4(w[j], tp.cls, tp.lower, nl, nd);
                L |= nl << (4 * j);
                D |= nd << (4 * j);
            }
            uint4* d = reinterpret_cast<uint4*>(ch8 + base);
            d[0] = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            d[1] = make_uint4(lw[4], lw[5], lw[6], lw[7]);
        } else {
            const uint4* p = reinterpret_cast<const uint4*>((const u16*)tv.text + x0);
            u32 sh[8];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint4 v = __ldg(p + q);
                const u32 ww[4] = {v.x, v.y, v.z, v.w};
                u32 o[4];
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    u32 k0, k1;
                    const u32 l0 = classify_unit(ww[r] & 0xFFFF, tp.cls, tp.lower, k0);
                    const u32 l1 = classify_unit(ww[r] >> 16, tp.cls, tp.lower, k1);
                    o[r] = l0 | (l1 << 16);
                    const int j = q * 8 + r * 2;
                    L |= ((k0 & 1u) << j) | ((k1 & 1u) << (j + 1));
                    D |= ((k0 >> 1) << j) | ((k1 >> 1) << (j + 1));
                    const u32 s0 = l0 < 128 ? l0 : 0xFFu, s1 = l1 < 128 ? l1 : 0xFFu;
                    const u32 pair = s0 | (s1 << 8);
                    if (r & 1) sh[q * 2 + (r >> 1)] |= pair << 16; else sh[q * 2 + (r >> 1)] = pair;
                }
                reinterpret_cast<uint4*>(ch16 + base)[q] = make_uint4(o[0], o[1], o[2], o[3]);
            }
            uint4* d = reinterpret_cast<uint4*>(ch8 + base);
            d[0] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
            d[1] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
        }
    } else {
        classify_span_scalar<ENC>(tv, tp, t0, base, ch8, ch16, L, D);
    }
    sm.mL[1 + c] = L;
    sm.mD[1 + c] = D;
    sm.mP[1 + c] = __ldg(tv.page_bits + (t0 >> 5) + c);
}

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

        // ---- 1. load, lower-case, classify: 32 units per thread -> one mask word per class ----------------
        // chunk c covers units [32c, 32c + 32) of the tile; chunks TK_THREADS and TK_THREADS + 1 are the look-ahead
        {
            classify_chunk<ENC>(tv, tp, t0, tid, ch8, ch16, sm);
            if (tid < TK_LA / TK_CPT) classify_chunk<ENC>(tv, tp, t0, TK_THREADS + tid, ch8, ch16, sm);
            if (tid == 32) {   // the unit before the tile
                u32 k = 0;
                if (t0 > 0) classify_unit(load_text_unit<ENC>(tv.text, t0 - 1), tp.cls, tp.lower, k);
                sm.mL[0] = (k & 1u) << 31;
                sm.mD[0] = (k >> 1) << 31;
            }
        }
        __syncthreads();

        // ---- 3. token starts and lengths by bit arithmetic ---------------------------------------------
        const u32 L = sm.mL[1 + tid], D = sm.mD[1 + tid], P = sm.mP[1 + tid];
        const u32 L1 = sm.mL[2 + tid], D1 = sm.mD[2 + tid], P1 = sm.mP[2 + tid];
        const u32 L2 = sm.mL[3 + tid], D2 = sm.mD[3 + tid], P2 = sm.mP[3 + tid];
        const u32 C = L | D, C1 = L1 | D1;
        const u32 S = (L & ~((L << 1) | (sm.mL[tid] >> 31))) | (D & ~((D << 1) | (sm.mD[tid] >> 31))) | (C & P);
        const u32 S1 = (L1 & ~((L1 << 1) | (L >> 31))) | (D1 & ~((D1 << 1) | (D >> 31))) | (C1 & P1);
        // continuation bits: units that extend the run started to their left
        const u64 rl = ((u64)(L1 & ~S1) << 32) | (u64)(L & ~S), rd = ((u64)(D1 & ~S1) << 32) | (u64)(D & ~S);
        u32 VL = S & L, VD = S & D;   // starts of runs with at least min_len units (Build.cs:528, lower bound)
        for (int j = 1; j < tp.min_len && j <= 32; j++) { VL &= (u32)(rl >> j); VD &= (u32)(rd >> j); }
        const u32 V = VL | VD;
        // break masks: a unit that does not continue a letter (digit) run, or starts a page
        const u64 bl01 = ((u64)(~L1 | P1) << 32) | (u64)(~L | P), bd01 = ((u64)(~D1 | P1) << 32) | (u64)(~D | P);
        const u64 bl2 = (u64)(~L2 | P2), bd2 = (u64)(~D2 | P2);
        u32 ntok;
        u32 off = block_excl_scan<u32, TK_THREADS>((u32)__popc(V), sm.scan, &ntok);
        for (u32 s = V; s; s &= s - 1) {
            const int i = __ffs(s) - 1;
            const bool dig = (D >> i) & 1u;
            const u64 x = ((dig ? bd01 : bl01) >> (i + 1)) | ((dig ? bd2 : bl2) << (63 - i));
            int len = x ? __ffsll((long long)x) : 99;
            if (len > tp.max_len || len < tp.min_len) len = 0;   // longer than max_len: the whole match is skipped (Build.cs:528); record dropped below
            tokl[off++] = (u32)(tid * TK_CPT + i) | ((u32)len << 13) | ((dig ? 1u : 0u) << 19);
        }
        __syncthreads();

        // ---- 4. per token: exact key, one probe, token record --------------------------------------------
        const i64 tdelta = tv.tile_delta[tile];
        u64* out = tp.tok + (u64)tile * tp.tok_cap;
        for (u32 k = tid; k < ntok; k += TK_THREADS) {
            const u32 e = tokl[k];
            const int start = (int)(e & 0x1FFF), len = (int)((e >> 13) & 63);
            const bool dig = (e >> 19) & 1u;
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
                const u32* w32 = reinterpret_cast<const u32*>(ch8 + (start & ~3));
                const int shb = (start & 3) * 8;
                auto chunk = [&](int j) -> u32 {
                    u32 v = __funnelshift_r(w32[j], w32[j + 1], shb);
                    const int rem = len - 4 * j;
                    if (rem < 4) v &= (1u << (8 * rem)) - 1u;
                    return v;
                };
                u64 lo = 0, hi = 0;
                u32 hb = 0;
                bool exact;
                if (!dig) {
                    u32 v = chunk(0); hb |= v; lo = pack_alpha4(v);
                    if (len > 4) { v = chunk(1); hb |= v; lo |= (u64)pack_alpha4(v) << 20; }
                    if (len > 8) { v = chunk(2); hb |= v; lo |= (u64)pack_alpha4(v) << 40; }
                    if (len > 12) { v = chunk(3); hb |= v; const u64 u = pack_alpha4(v); lo |= u << 60; hi = u >> 4; }
                    if (len > 16) { v = chunk(4); hb |= v; hi |= (u64)pack_alpha4(v) << 16; }
                    if (len > 20) { v = chunk(5); hb |= v; hi |= (u64)pack_alpha4(v) << 36; }
                    if (len > 24) { v = chunk(6); hb |= v; hi |= (u64)pack_alpha4(v) << 56; }
                    hi |= (u64)WK_ALPHA << WK_TAG_SHIFT;
                    exact = len <= 25 && !(hb & 0x80808080u);
                } else {
                    u32 v = chunk(0); hb |= v; lo = pack_digit4(v);
                    if (len > 4) { v = chunk(1); hb |= v; lo |= (u64)pack_digit4(v) << 16; }
                    if (len > 8) { v = chunk(2); hb |= v; lo |= (u64)pack_digit4(v) << 32; }
                    if (len > 12) { v = chunk(3); hb |= v; lo |= (u64)pack_digit4(v) << 48; }
                    if (len > 16) { v = chunk(4); hb |= v; hi |= (u64)pack_digit4(v); }
                    if (len > 20) { v = chunk(5); hb |= v; hi |= (u64)pack_digit4(v) << 16; }
                    if (len > 24) { v = chunk(6); hb |= v; hi |= (u64)pack_digit4(v) << 32; }
                    if (len > 28) { v = chunk(7); hb |= v; hi |= (u64)pack_digit4(v) << 48; }
                    hi |= (u64)WK_DIGIT << WK_TAG_SHIFT;
                    exact = len <= 30 && !(hb & 0x80808080u);
                }
                if (exact) slot = wordtab_intern(tp.tab, lo, hi);
                else if (ENC == 1) slot = wordtab_intern_units(tp.tab, (const u8*)(ch8 + start), len);
                else slot = wordtab_intern_units(tp.tab, (const u16*)(ch16 + start), len);
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
// lane-level classification of 32 units at unit offset x0: lower-cased bytes to ch8 (units to ch16), masks out
template <int ENC>
DCD_D void classify32(const void* text, u64 n_units, u64 x0, const u8* __restrict__ cls, const u16* __restrict__ lower, u8* ch8,
                      u16* ch16, u32& L, u32& D) {
    L = D = 0;
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
                    u32 k0, k1;
                    const u32 l0 = classify_unit(ww[r] & 0xFFFF, cls, lower, k0);
                    const u32 l1 = classify_unit(ww[r] >> 16, cls, lower, k1);
                    o[r] = l0 | (l1 << 16);
                    const int j = q * 8 + r * 2;
                    L |= ((k0 & 1u) << j) | ((k1 & 1u) << (j + 1));
                    D |= ((k0 >> 1) << j) | ((k1 >> 1) << (j + 1));
                    const u32 s0 = l0 < 128 ? l0 : 0xFFu, s1 = l1 < 128 ? l1 : 0xFFu;
                    const u32 pair = s0 | (s1 << 8);
                    if (r & 1) sh[q * 2 + (r >> 1)] |= pair << 16; else sh[q * 2 + (r >> 1)] = pair;
                }
                reinterpret_cast<uint4*>(ch16)[q] = make_uint4(o[0], o[1], o[2], o[3]);
            }
            uint4* d = reinterpret_cast<uint4*>(ch8);
            d[0] = make_uint4(sh[0], sh[1], sh[2], sh[3]);
            d[1] = make_uint4(sh[4], sh[5], sh[6], sh[7]);
        }
    } else {
        for (int j = 0; j < 32; j++) {
            const u64 x = x0 + (u64)j;
            u32 l = 0, k = 0;
            if (x < n_units) l = classify_unit(load_text_unit<ENC>(text, x), cls, lower, k);
            if (ENC == 1) ch8[j] = (u8)l;
            else { ch16[j] = (u16)l; ch8[j] = l < 128 ? (u8)l : (u8)0xFF; }
            L |= (k & 1u) << j;
            D |= (k >> 1) << j;
        }
    }
}

constexpr int T2_CH8 = T2_WIN + 64;                    // window + over-read pad of the key builder
template <int ENC> struct Tok2WarpSmem {
    u8 ch8[T2_CH8];
    u16 ch16[ENC == 2 ? T2_WIN : 8];
    u32 tokl[T2_CAP + 4];
};

// exact 128-bit key of the token at ch8[start, start + len), or a table insert by units when it is not [a-z]+ / [0-9]+
template <int ENC>
DCD_D u32 intern_token(const WordTable& tab, const u8* ch8, const u16* ch16, int start, int len, bool dig) {
    const u32* w32 = reinterpret_cast<const u32*>(ch8 + (start & ~3));
    const int shb = (start & 3) * 8;
    auto chunk = [&](int j) -> u32 {
        u32 v = __funnelshift_r(w32[j], w32[j + 1], shb);
        const int rem = len - 4 * j;
        if (rem < 4) v &= (1u << (8 * rem)) - 1u;
        return v;
    };
    u64 lo = 0, hi = 0;
    u32 hb = 0;
    bool exact;
    if (!dig) {
        u32 v = chunk(0); hb |= v; lo = pack_alpha4(v);
        if (len > 4) { v = chunk(1); hb |= v; lo |= (u64)pack_alpha4(v) << 20; }
        if (len > 8) { v = chunk(2); hb |= v; lo |= (u64)pack_alpha4(v) << 40; }
        if (len > 12) { v = chunk(3); hb |= v; const u64 u = pack_alpha4(v); lo |= u << 60; hi = u >> 4; }
        if (len > 16) { v = chunk(4); hb |= v; hi |= (u64)pack_alpha4(v) << 16; }
        if (len > 20) { v = chunk(5); hb |= v; hi |= (u64)pack_alpha4(v) << 36; }
        if (len > 24) { v = chunk(6); hb |= v; hi |= (u64)pack_alpha4(v) << 56; }
        hi |= (u64)WK_ALPHA << WK_TAG_SHIFT;
        exact = len <= 25 && !(hb & 0x80808080u);
    } else {
        u32 v = chunk(0); hb |= v; lo = pack_digit4(v);
        if (len > 4) { v = chunk(1); hb |= v; lo |= (u64)pack_digit4(v) << 16; }
        if (len > 8) { v = chunk(2); hb |= v; lo |= (u64)pack_digit4(v) << 32; }
        if (len > 12) { v = chunk(3); hb |= v; lo |= (u64)pack_digit4(v) << 48; }
        if (len > 16) { v = chunk(4); hb |= v; hi |= (u64)pack_digit4(v); }
        if (len > 20) { v = chunk(5); hb |= v; hi |= (u64)pack_digit4(v) << 16; }
        if (len > 24) { v = chunk(6); hb |= v; hi |= (u64)pack_digit4(v) << 32; }
        if (len > 28) { v = chunk(7); hb |= v; hi |= (u64)pack_digit4(v) << 48; }
        hi |= (u64)WK_DIGIT << WK_TAG_SHIFT;
        exact = len <= 30 && !(hb & 0x80808080u);
    }
    if (exact) return wordtab_intern(tab, lo, hi);
    if (ENC == 1) return wordtab_intern_units(tab, (const u8*)(ch8 + start), len);
    return wordtab_intern_units(tab, (const u16*)(ch16 + start), len);
}

template <int ENC>
__global__ void __launch_bounds__(T2_WARPS * 32, 8) k_tokenize2(Tok2Params tp) {
    extern __shared__ __align__(16) unsigned char smraw[];
    Tok2WarpSmem<ENC>& sm = reinterpret_cast<Tok2WarpSmem<ENC>*>(smraw)[threadIdx.x >> 5];
    const u32 lane = lane_id();
    const u32 gw = blockIdx.x * T2_WARPS + (threadIdx.x >> 5), nw = gridDim.x * T2_WARPS;
    for (int i = T2_WIN + (int)lane; i < T2_CH8; i += 32) sm.ch8[i] = 0;
    u64 my_tokens = 0;
    for (u32 tile = tp.tile_begin + gw; tile < tp.ntiles; tile += nw) {
        const u64 t0 = (u64)tile * T2_OWN;
        __syncwarp();   // the previous tile's token phase is done with ch8 / tokl
        u32 L, D;
        classify32<ENC>(tp.text, tp.n_units, t0 + 32u * lane, tp.cls, tp.lower, sm.ch8 + 32 * lane, sm.ch16 + 32 * lane, L, D);
        const u32 P = __ldg(tp.page_bits + (t0 >> 5) + lane);
        // class of the unit before this lane's 32 (lane 0: the unit before the tile)
        u32 pl = __shfl_up_sync(0xffffffffu, L, 1) >> 31, pd = __shfl_up_sync(0xffffffffu, D, 1) >> 31;
        if (lane == 0) {
            u32 k = 0;
            if (t0 > 0) classify_unit(load_text_unit<ENC>(tp.text, t0 - 1), tp.cls, tp.lower, k);
            pl = k & 1u;
            pd = k >> 1;
        }
        u32 L1 = __shfl_down_sync(0xffffffffu, L, 1), D1 = __shfl_down_sync(0xffffffffu, D, 1), P1 = __shfl_down_sync(0xffffffffu, P, 1);
        u32 L2 = __shfl_down_sync(0xffffffffu, L, 2), D2 = __shfl_down_sync(0xffffffffu, D, 2), P2 = __shfl_down_sync(0xffffffffu, P, 2);
        if (lane >= 31) { L1 = D1 = P1 = 0; }
        if (lane >= 30) { L2 = D2 = P2 = 0; }
        const u32 C = L | D, C1 = L1 | D1;
        const u32 S = (L & ~((L << 1) | pl)) | (D & ~((D << 1) | pd)) | (C & P);
        const u32 S1 = (L1 & ~((L1 << 1) | (L >> 31))) | (D1 & ~((D1 << 1) | (D >> 31))) | (C1 & P1);
        const u64 rl = ((u64)(L1 & ~S1) << 32) | (u64)(L & ~S), rd = ((u64)(D1 & ~S1) << 32) | (u64)(D & ~S);
        u32 VL = S & L, VD = S & D;   // starts of runs with at least min_len units (Build.cs:528, lower bound)
        for (int j = 1; j < tp.min_len && j <= 32; j++) { VL &= (u32)(rl >> j); VD &= (u32)(rd >> j); }
        const u32 V = lane < T2_OWN / 32 ? (VL | VD) : 0u;
        const u64 bl01 = ((u64)(~L1 | P1) << 32) | (u64)(~L | P), bd01 = ((u64)(~D1 | P1) << 32) | (u64)(~D | P);
        const u64 bl2 = (u64)(~L2 | P2), bd2 = (u64)(~D2 | P2);
        const u32 cnt = (u32)__popc(V);
        const u32 incl = warp_incl_scan(cnt);
        const u32 ntok = __shfl_sync(0xffffffffu, incl, 31);
        u32 off = incl - cnt;
        for (u32 sbits = V; sbits; sbits &= sbits - 1) {
            const int i = __ffs(sbits) - 1;
            const bool dig = (D >> i) & 1u;
            const u64 x = ((dig ? bd01 : bl01) >> (i + 1)) | ((dig ? bd2 : bl2) << (63 - i));
            int len = x ? __ffsll((long long)x) : 99;
            if (len > tp.max_len || len < tp.min_len) len = 0;   // longer than max_len: the whole match is skipped (Build.cs:528)
            sm.tokl[off++] = (u32)(lane * 32 + i) | ((u32)len << 10) | ((dig ? 1u : 0u) << 16);
        }
        __syncwarp();
        u32* out = tp.rec + (u64)tile * T2_CAP;
        for (u32 k = lane; k < ntok; k += 32) {
            const u32 e = sm.tokl[k];
            const int start = (int)(e & 0x3FF), len = (int)((e >> 10) & 63);
            u32 slot = T2_SLOT_NONE;
            if (len) {
                slot = intern_token<ENC>(tp.tab, sm.ch8, sm.ch16, start, len, (e >> 16) & 1u);
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
