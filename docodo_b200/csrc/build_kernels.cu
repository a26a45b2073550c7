// K1-K4: tokenizer, distinct-word table, GetWordCodes, posting emit.  Hand-written for sm_100a.
//
// Replaces the per-token loop of IndexTask (Build.cs:477-536): ToLower (482), Regex \p{L}+|\p{N}+ (526), the
// 3..32 length filter (528), Builder.AddWord -> GetWordCodes (302-316, 175-247) and Builder.Add (324-341).
//
// Data flow (two passes over the text, both HBM-streaming with 128-bit loads into a shared-memory tile):
//   pass 1  k_tokenize<INSERT>: tile -> lower-case + class in smem -> token list -> 64-bit hash -> find-or-insert
//           into the device string table (distinct words; L2-resident hot set).
//   K4      k_word_codes: one thread per DISTINCT word: stop flag, Snowball stem, vocabulary probe -> up to 6 codes.
//   (host)  ordinal ranking of the code strings (radix sort, sort.cu) -> code ids.
//   pass 2  k_tokenize<EMIT>: same tile walk, table lookup, packed keys (codeId<<40 | coord) written in coordinate
//           order through a decoupled look-back, with the radix-sort digit histograms fused in.
#include "build_kernels.cuh"
#include "stem_en.cuh"
#include "stem_ru.cuh"

namespace dcd {

// ---------------------------------------------------------------------------------------------------
// tile_first_page: page containing the first unit of each tile (binary search, one thread per tile)
// ---------------------------------------------------------------------------------------------------
__global__ void k_tile_first_page(const u64* __restrict__ page_off, u64 npages, u32* __restrict__ out, u32 ntiles) {
    u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    u64 x = (u64)t * TOK_TILE;
    // last p with page_off[p] <= x  (page_off[0] == 0 <= x always)
    u64 lo = 0, hi = npages;   // invariant: page_off[lo] <= x, page_off[hi] > x or hi == npages
    while (hi - lo > 1) {
        u64 mid = (lo + hi) >> 1;
        if (page_off[mid] <= x) lo = mid; else hi = mid;
    }
    out[t] = (u32)lo;
}
void launch_tile_first_page(const u64* page_off, u64 npages, u32* out, u32 ntiles, cudaStream_t s) {
    if (!ntiles) return;
    k_tile_first_page<<<(ntiles + 255) / 256, 256, 0, s>>>(page_off, npages, out, ntiles);
}

// ---------------------------------------------------------------------------------------------------
// tokenizer
// ---------------------------------------------------------------------------------------------------
struct TokSmem {
    u16 ch[TOK_TILE + TOK_LA + 8];     // [0] = unit before the tile, [1..TILE] tile, then look-ahead (lower-cased)
    u8 cl[TOK_TILE + TOK_LA + 8];      // class (0 none, 1 letter, 2 number) | 0x80 page starts here | 0x40 beyond text end
    u32 tok[TOK_MAX_TOKENS];           // (local index << 8) | length
    u32 tok_entry[TOK_MAX_TOKENS];     // EMIT: entry id of the token's word
    u64 pg_off[TOK_PG_CACHE];
    u32 scan[TOK_THREADS / 32 + 1];
    u64 scan64[TOK_THREADS / 32 + 1];
    u32 hist[4][256];
    u32 ntok;
    u32 npg;
    u32 tile;
    u64 tile_base;
};

DCD_D void classify_store(TokSmem& sm, int idx, u32 c, const u8* __restrict__ cls, const u16* __restrict__ lower) {
    u32 l, k;
    if (c < 128) {
        bool up = c >= 'A' && c <= 'Z';
        l = up ? c + 32 : c;
        k = ((l >= 'a' && l <= 'z') ? 1u : 0u) | ((c >= '0' && c <= '9') ? 2u : 0u);
    } else {
        l = __ldg(lower + c);
        u32 b = __ldg(cls + l);
        k = (b & CLS_L) ? 1u : ((b & CLS_N) ? 2u : 0u);
    }
    sm.ch[idx] = (u16)l;
    sm.cl[idx] = (u8)k;
}

DCD_D u32 load_unit(const TextView& tv, u64 x) {
    return tv.enc == 1 ? (u32)((const u8*)tv.text)[x] : (u32)((const u16*)tv.text)[x];
}

template <int MODE>   // 0 = INSERT (pass 1), 1 = EMIT (pass 2)
__global__ void __launch_bounds__(TOK_THREADS) k_tokenize(TextView tv, BuildParams bp, EmitParams ep, u32* ticket,
                                                           u64* token_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TokSmem& sm = *reinterpret_cast<TokSmem*>(smem_raw);
    const int tid = threadIdx.x;
    u64 my_tokens = 0;
    if (MODE == 1) {
        for (int i = tid; i < 4 * 256; i += TOK_THREADS) (&sm.hist[0][0])[i] = 0;
    }
    while (true) {
        __syncthreads();
        if (tid == 0) sm.tile = atomicAdd(ticket, 1u);
        __syncthreads();
        const u32 tile = sm.tile;
        if (tile >= tv.ntiles) break;
        const u64 t0 = (u64)tile * TOK_TILE;
        const u64 remaining = tv.n_units - t0;
        const int tile_n = remaining < (u64)TOK_TILE ? (int)remaining : TOK_TILE;

        // ---- 1. load + lower-case + classify -------------------------------------------------
        {
            const int base = tid * TOK_PER_THREAD;
            if (base + TOK_PER_THREAD <= tile_n) {
                if (tv.enc == 1) {
                    uint4 v = ld_stream_u128((const u8*)tv.text + t0 + base);
                    u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int k = 0; k < 16; k++) classify_store(sm, 1 + base + k, (w[k >> 2] >> (8 * (k & 3))) & 0xFF, bp.cls, bp.lower);
                } else {
                    const u16* p = (const u16*)tv.text + t0 + base;
                    uint4 v0 = ld_stream_u128(p), v1 = ld_stream_u128(p + 8);
                    u32 w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                    for (int k = 0; k < 16; k++) classify_store(sm, 1 + base + k, (w[k >> 1] >> (16 * (k & 1))) & 0xFFFF, bp.cls, bp.lower);
                }
            } else {
                for (int k = 0; k < TOK_PER_THREAD; k++) {
                    int i = base + k;
                    if (i < tile_n) classify_store(sm, 1 + i, load_unit(tv, t0 + i), bp.cls, bp.lower);
                    else { sm.ch[1 + i] = 0; sm.cl[1 + i] = 0x40; }
                }
            }
            // look-behind unit and look-ahead units
            if (tid == 0) {
                if (t0 > 0) classify_store(sm, 0, load_unit(tv, t0 - 1), bp.cls, bp.lower);
                else { sm.ch[0] = 0; sm.cl[0] = 0; }
            }
            if (tid >= 32 && tid < 32 + TOK_LA) {
                int i = TOK_TILE + (tid - 32);
                u64 x = t0 + i;
                if (x < tv.n_units) classify_store(sm, 1 + i, load_unit(tv, x), bp.cls, bp.lower);
                else { sm.ch[1 + i] = 0; sm.cl[1 + i] = 0x40; }
            }
            if (tid == 0) { sm.ntok = 0; sm.npg = 0; }
        }
        __syncthreads();

        // ---- 2. page starts inside [t0, t0 + TILE + LA) ----------------------------------------
        const u32 p0 = tv.tile_first_page[tile];
        {
            const u64 limit = t0 + TOK_TILE + TOK_LA;
            for (u64 p = (u64)p0 + tid; p < tv.npages; p += TOK_THREADS) {
                u64 off = tv.page_off[p];
                if (off >= limit) break;
                if (off >= t0) sm.cl[1 + (int)(off - t0)] |= 0x80;   // byte stores to distinct or identical-valued addresses
                u32 slot = (u32)(p - p0);
                if (slot < TOK_PG_CACHE) sm.pg_off[slot] = off;
            }
            if (tid == 0) {
                // number of pages whose start lies before the end of the tile (upper bound for the cache walk)
                u64 p = p0;
                u32 cnt = 0;
                while (p < tv.npages && cnt < TOK_PG_CACHE && tv.page_off[p] < t0 + TOK_TILE) { p++; cnt++; }
                sm.npg = (p < tv.npages && tv.page_off[p] < t0 + TOK_TILE) ? 0xFFFFFFFFu : cnt;   // overflow -> global search
            }
        }
        __syncthreads();

        // ---- 3. token starts owned by this thread ---------------------------------------------
        // A token starts at i when class != 0 and (previous class differs, or a page starts at i).
        auto token_len = [&](int i) -> int {   // i = local index into the tile; returns 0 when not a valid token
            const u8 c = sm.cl[1 + i];
            const u32 k = c & 3u;
            if (k == 0) return 0;
            if (i > 0 || t0 > 0) {
                const u8 pc = sm.cl[i];
                if ((pc & 3u) == k && !(c & 0x80)) return 0;
            }
            int j = i + 1;
            const int lim = i + bp.max_len + 1;
            while (j < lim) {
                const u8 d = sm.cl[1 + j];
                if ((d & 3u) != k || (d & 0xC0)) break;
                j++;
            }
            if (j == lim) return 0;   // longer than max_len: Regex match is skipped as a whole (Build.cs:528)
            const int len = j - i;
            return len >= bp.min_len ? len : 0;
        };
        u32 cnt = 0;
        u32 lens[TOK_PER_THREAD];
#pragma unroll
        for (int k = 0; k < TOK_PER_THREAD; k++) {
            int i = tid * TOK_PER_THREAD + k;
            int l = i < tile_n ? token_len(i) : 0;
            lens[k] = (u32)l;
            cnt += l ? 1u : 0u;
        }
        u32 total_tok;
        u32 off = block_excl_scan<u32, TOK_THREADS>(cnt, sm.scan, &total_tok);
#pragma unroll
        for (int k = 0; k < TOK_PER_THREAD; k++)
            if (lens[k]) sm.tok[off++] = ((u32)(tid * TOK_PER_THREAD + k) << 8) | lens[k];
        __syncthreads();
        const int ntok = (int)total_tok;

        // page of a unit offset x (absolute): last page with page_off <= x
        auto page_of = [&](u64 x) -> u64 {
            if (sm.npg != 0xFFFFFFFFu) {
                u32 lo = 0, hi = sm.npg;   // pg_off[lo] <= x
                while (hi - lo > 1) {
                    u32 mid = (lo + hi) >> 1;
                    if (sm.pg_off[mid] <= x) lo = mid; else hi = mid;
                }
                return (u64)p0 + lo;
            }
            u64 lo = p0, hi = tv.npages;
            while (hi - lo > 1) {
                u64 mid = (lo + hi) >> 1;
                if (tv.page_off[mid] <= x) lo = mid; else hi = mid;
            }
            return lo;
        };

        // ---- 4. per token: hash, table ---------------------------------------------------------
        if (MODE == 0) {
            for (int k = tid; k < ntok; k += TOK_THREADS) {
                const u32 t = sm.tok[k];
                const int i = (int)(t >> 8), len = (int)(t & 0xFF);
                const u64 pg = page_of(t0 + i);
                if (tv.page_delta[pg] == PAGE_SKIP) continue;
                const u16* w = &sm.ch[1 + i];
                const u64 h = hash_units(w, len);
                u32 e = strtab_find(bp.tab, w, len, h);            // L1-cached fast path (entries are write-once)
                if (e == ENT_NONE || !(bp.tab.ent_flags[e] & EF_WORD)) e = strtab_insert(bp.tab, w, len, h, EF_WORD);
                my_tokens++;
            }
        } else {
            // 4a. lookup; n codes per token
            u32 mycodes = 0;
            const int per = (ntok + TOK_THREADS - 1) / TOK_THREADS;
            const int k0 = tid * per, k1 = min(ntok, k0 + per);
            for (int k = k0; k < k1; k++) {
                const u32 t = sm.tok[k];
                const int i = (int)(t >> 8), len = (int)(t & 0xFF);
                const u64 pg = page_of(t0 + i);
                u32 e = ENT_NONE;
                if (tv.page_delta[pg] != PAGE_SKIP) {
                    const u16* w = &sm.ch[1 + i];
                    e = strtab_find(bp.tab, w, len, hash_units(w, len));
                    if (e == ENT_NONE) atomicExch(ep.error + 1, 1u);
                    else mycodes += ep.word_codes[(u64)e * 8];
                    my_tokens++;
                }
                sm.tok_entry[k] = e;
            }
            u32 tile_codes;
            u32 myoff = block_excl_scan<u32, TOK_THREADS>(mycodes, sm.scan, &tile_codes);
            if (warp_id() == 0) {
                u64 ex = lookback_warp(ep.lookback, (int)tile, (u64)tile_codes);
                if (lane_id() == 0) sm.tile_base = ex;
            }
            __syncthreads();
            u64 dst = sm.tile_base + myoff;
            if (sm.tile_base + tile_codes > ep.capacity) {
                if (tid == 0) atomicExch(ep.error + 0, 1u);
            } else {
                for (int k = k0; k < k1; k++) {
                    const u32 e = sm.tok_entry[k];
                    if (e == ENT_NONE) continue;
                    const u32 t = sm.tok[k];
                    const u64 x = t0 + (t >> 8);
                    const u64 pg = page_of(x);
                    const u64 coord = (u64)((i64)x + tv.page_delta[pg]);
                    if (coord > COORD_MASK) atomicExch(ep.error + 2, 1u);
                    const u32* wc = ep.word_codes + (u64)e * 8;
                    const u32 n = wc[0];
                    for (u32 c = 0; c < n; c++) {
                        const u32 id = wc[1 + c];
                        ep.keys[dst++] = ((u64)id << COORD_BITS) | (coord & COORD_MASK);
                        for (int ps = 0; ps < ep.npasses; ps++)
                            atomicAdd(&sm.hist[ps][(id >> ep.pass_shift[ps]) & ((1u << ep.pass_bits[ps]) - 1)], 1u);
                    }
                }
            }
            if (tile == tv.ntiles - 1 && tid == 0) ep.total[0] = sm.tile_base + tile_codes;
        }
    }
    // ---- epilogue: flush per-CTA counters -----------------------------------------------------------
    my_tokens = warp_sum(my_tokens);
    if (lane_id() == 0 && my_tokens) atomicAdd(token_count, my_tokens);
    if (MODE == 1) {
        __syncthreads();
        for (int i = tid; i < ep.npasses * 256; i += TOK_THREADS) {
            u32 v = (&sm.hist[0][0])[i];
            if (v) atomicAdd(ep.hist + i, (u64)v);
        }
    }
}

static int tok_grid(int nsm, const void* fn) {
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, TOK_THREADS, sizeof(TokSmem));
    if (occ < 1) occ = 1;
    return nsm * occ;
}

void launch_tokenize_insert(const TextView& tv, const BuildParams& bp, u32* ticket, u64* token_count, int nsm, cudaStream_t s) {
    if (!tv.ntiles) return;
    auto fn = k_tokenize<0>;
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TokSmem));
    EmitParams ep{};
    fn<<<tok_grid(nsm, (const void*)fn), TOK_THREADS, sizeof(TokSmem), s>>>(tv, bp, ep, ticket, token_count);
}
void launch_tokenize_emit(const TextView& tv, const BuildParams& bp, const EmitParams& ep, u32* ticket, int nsm, cudaStream_t s) {
    if (!tv.ntiles) return;
    auto fn = k_tokenize<1>;
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TokSmem));
    fn<<<tok_grid(nsm, (const void*)fn), TOK_THREADS, sizeof(TokSmem), s>>>(tv, bp, ep, ticket, ep.total + 1);
}

// ---------------------------------------------------------------------------------------------------
// K4: GetWordCodes per distinct word (Build.cs:175-247)
// ---------------------------------------------------------------------------------------------------
DCD_D bool in_range_all(const u16* w, int n, int kind) {   // !Regex.IsMatch(word, "[^range]"), Build.cs:216; Index.cs:175-181
    for (int i = 0; i < n; i++) {
        const u16 c = w[i];
        bool ok;
        switch (kind) {
            case 0: ok = c >= '0' && c <= '9'; break;
            case 1: ok = (c >= 0x0430 && c <= 0x044F) || c == 0x0451; break;                       // а-яё
            case 2: ok = c >= 'a' && c <= 'z'; break;
            case 3: ok = (c >= 'a' && c <= 'z') || c == 0x1E9E || c == 0xE4 || c == 0xFC || c == 0xF6; break;   // a-zẞäüö
            default:
                ok = (c >= 'a' && c <= 'z') || c == 0xE9 || c == 0xE2 || c == 0xE0 || c == 0xEA || c == 0xE8 || c == 0xEB ||
                     c == 0xE7 || c == 0xEE || c == 0xEF || c == 0xF4 || c == 0xFB || c == 0xF9 || c == 0xFC || c == 0xFF;
        }
        if (!ok) return false;
    }
    return true;
}

DCD_D int run_stemmer(int kind, u16* s, int n) {
    if (kind == 1) return en::stem(s, n);
    if (kind == 2) return ru::stem(s, n);
    return n;
}

__global__ void __launch_bounds__(128) k_word_codes(BuildParams bp, u32 n_entries, u32* __restrict__ word_codes) {
    const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_entries) return;
    const StrTable& t = bp.tab;
    u32* out = word_codes + (u64)e * 8;
    const u32 flags = t.ent_flags[e];
    if (!(flags & EF_WORD)) { out[0] = 0; return; }
    const int len = (int)t.ent_len[e];
    const u16* src = t.arena + t.ent_off[e];
    u16 w[MAX_KEY_UNITS + 2];
    for (int i = 0; i < len; i++) w[i] = src[i];
    u32 n = 0;
    if (len == 0) { out[0] = 0; return; }                      // Build.cs:178
    if (w[0] >= '0' && w[0] <= '9') {                          // Build.cs:180-181
        atomicOr(t.ent_flags + e, EF_CODE);
        out[0] = 1; out[1] = e;
        return;
    }
    if (flags & EF_STOP) { out[0] = 0; return; }               // Build.cs:185
    atomicOr(t.ent_flags + e, EF_CODE);                        // Build.cs:188
    out[1 + n++] = e;
    u16 st[MAX_KEY_UNITS + 2];                                 // `stemmed`
    int ls = len;
    for (int i = 0; i < len; i++) st[i] = w[i];
    u16 first[MAX_KEY_UNITS + 2];                              // `firststemmed`
    int lf = 0;
    int nG = 0;
    for (int v = 0; v < bp.nvocs; v++) {                       // Build.cs:192-204
        const VocabDev& voc = bp.vocs[v];
        if (w[0] >= voc.begin && w[0] <= voc.end) {
            for (int i = 0; i < len; i++) st[i] = w[i];
            ls = run_stemmer(voc.stemmer, st, len);
            const u32 ve = strtab_find(voc.tab, st, ls, hash_units(st, ls));
            nG = ve == ENT_NONE ? 0 : (int)voc.tab.ent_value[ve];
            if (nG != 0) {
                // "#" + String.Format("{0:X}", (nVoc << 24) | (nG & 0xFFFFFF)), Index.cs:196
                u32 val = ((u32)v << 24) | ((u32)nG & 0xFFFFFFu);
                u16 code[10];
                int lc = 0;
                code[lc++] = '#';
                int digits = 1;
                while (digits < 8 && (val >> (4 * digits)) != 0) digits++;
                for (int d = digits - 1; d >= 0; d--) {
                    u32 x = (val >> (4 * d)) & 0xF;
                    code[lc++] = (u16)(x < 10 ? '0' + x : 'A' + x - 10);
                }
                u32 ce = strtab_insert(t, code, lc, hash_units(code, lc), EF_CODE);
                if (ce != ENT_NONE && n < MAX_CODES) out[1 + n++] = ce;
            }
        }
        if (lf == 0) { lf = ls; for (int i = 0; i < ls; i++) first[i] = st[i]; }
    }
    if (nG == 0) {                                             // Build.cs:206-238
        ls = lf;
        for (int i = 0; i < lf; i++) st[i] = first[i];
        if (bp.nvocs == 0) {
            for (int k = 0; k < 5; k++) {
                if (!(bp.stemmers_mask & (1 << k))) continue;
                if (in_range_all(w, len, k)) {
                    if (k == 1 || k == 2) {
                        for (int i = 0; i < len; i++) st[i] = w[i];
                        ls = run_stemmer(k == 1 ? 2 : 1, st, len);
                    }
                    break;
                }
            }
        }
        bool same = ls == len;
        if (same) for (int i = 0; i < len; i++) if (st[i] != w[i]) { same = false; break; }
        if (ls > 0 && !same) {
            u16 code[MAX_KEY_UNITS + 2];
            code[0] = '$';
            for (int i = 0; i < ls; i++) code[1 + i] = st[i];
            int lc = ls + 1;
            if (lc > MAX_KEY_UNITS) lc = MAX_KEY_UNITS;   // cannot happen: stems are never longer than the word (<= 47 after '$')
            u32 ce = strtab_insert(t, code, lc, hash_units(code, lc), EF_CODE);
            if (ce != ENT_NONE && n < MAX_CODES) out[1 + n++] = ce;
        }
    }
    out[0] = n;
}
void launch_word_codes(const BuildParams& bp, u32 n_entries, u32* word_codes, cudaStream_t s) {
    if (!n_entries) return;
    k_word_codes<<<(n_entries + 127) / 128, 128, 0, s>>>(bp, n_entries, word_codes);
}

__global__ void k_codes_to_ids(StrTable t, u32 n_entries, u32* __restrict__ word_codes) {
    const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_entries) return;
    u32* wc = word_codes + (u64)e * 8;
    const u32 n = wc[0];
    for (u32 c = 0; c < n && c < MAX_CODES; c++) wc[1 + c] = t.ent_value[wc[1 + c]];
}
void launch_codes_to_ids(const StrTable& tab, u32 n_entries, u32* word_codes, cudaStream_t s) {
    if (!n_entries) return;
    k_codes_to_ids<<<(n_entries + 255) / 256, 256, 0, s>>>(tab, n_entries, word_codes);
}

// ---------------------------------------------------------------------------------------------------
// host-supplied strings
// ---------------------------------------------------------------------------------------------------
__global__ void k_insert_strings(StrTable t, const u16* __restrict__ chars, const u32* __restrict__ offs, u32 n, u32 flags,
                                 const u32* __restrict__ values, u32* __restrict__ out_entries) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u16* s = chars + offs[i];
    const int len = (int)(offs[i + 1] - offs[i]);
    u32 e = strtab_insert(t, s, len, hash_units(s, len), flags);
    if (e != ENT_NONE && values) t.ent_value[e] = values[i];
    if (out_entries) out_entries[i] = e;
}
void launch_insert_strings(const StrTable& tab, const u16* chars, const u32* offs, u32 n, u32 flags, const u32* values,
                           u32* out_entries, cudaStream_t s) {
    if (!n) return;
    k_insert_strings<<<(n + 127) / 128, 128, 0, s>>>(tab, chars, offs, n, flags, values, out_entries);
}

__global__ void k_gather_flagged(StrTable t, u32 n_entries, u32 flag, u32* __restrict__ out, u32* __restrict__ out_n) {
    const u32 e = blockIdx.x * blockDim.x + threadIdx.x;
    const bool has = e < n_entries && (t.ent_flags[e] & flag);
    const u32 m = __ballot_sync(0xffffffffu, has);
    u32 base = 0;
    if (lane_id() == 0 && m) base = atomicAdd(out_n, (u32)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (has) out[base + __popc(m & ((1u << lane_id()) - 1))] = e;
}
void launch_gather_flagged(const StrTable& tab, u32 n_entries, u32 flag, u32* out, u32* out_n, cudaStream_t s) {
    if (!n_entries) return;
    k_gather_flagged<<<(n_entries + 255) / 256, 256, 0, s>>>(tab, n_entries, flag, out, out_n);
}

__global__ void k_emit_extras(StrTable t, const u32* __restrict__ word_codes, const u32* __restrict__ entries,
                              const u64* __restrict__ coords, const u8* __restrict__ kinds, u32 n, u64* __restrict__ keys,
                              u32* __restrict__ nkeys) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 e = entries[i];
    if (e == ENT_NONE) return;
    const u64 coord = coords[i] & COORD_MASK;
    if (kinds[i] == 1) {
        keys[atomicAdd(nkeys, 1u)] = ((u64)t.ent_value[e] << COORD_BITS) | coord;
    } else {
        const u32* wc = word_codes + (u64)e * 8;
        const u32 k = wc[0];
        if (!k) return;
        u32 at = atomicAdd(nkeys, k);
        for (u32 c = 0; c < k; c++) keys[at + c] = ((u64)wc[1 + c] << COORD_BITS) | coord;
    }
}
void launch_emit_extras(const StrTable& tab, const u32* word_codes, const u32* entries, const u64* coords, const u8* kinds,
                        u32 n, u64* keys, u32* nkeys, cudaStream_t s) {
    if (!n) return;
    k_emit_extras<<<(n + 127) / 128, 128, 0, s>>>(tab, word_codes, entries, coords, kinds, n, keys, nkeys);
}

// ---------------------------------------------------------------------------------------------------
// pages
// ---------------------------------------------------------------------------------------------------
__global__ void k_page_adv(const u64* __restrict__ page_off, u64 npages, u64* __restrict__ adv) {
    u64 p = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < npages) adv[p] = page_off[p + 1] - page_off[p];
}
void launch_page_adv(const u64* page_off, u64 npages, u64* adv, cudaStream_t s) {
    if (!npages) return;
    k_page_adv<<<(unsigned)((npages + 255) / 256), 256, 0, s>>>(page_off, npages, adv);
}
__global__ void k_page_delta(const u64* __restrict__ page_off, const u64* __restrict__ adv, const u64* __restrict__ coord_end,
                             const u8* __restrict__ kind, u64 npages, i64* __restrict__ delta) {
    u64 p = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npages) return;
    if (kind && kind[p]) delta[p] = PAGE_SKIP;
    else delta[p] = (i64)(coord_end[p] - adv[p]) - (i64)page_off[p];
}
void launch_page_delta(const u64* page_off, const u64* adv, const u64* coord_end, const u8* kind, u64 npages, i64* delta,
                       cudaStream_t s) {
    if (!npages) return;
    k_page_delta<<<(unsigned)((npages + 255) / 256), 256, 0, s>>>(page_off, adv, coord_end, kind, npages, delta);
}

}  // namespace dcd
