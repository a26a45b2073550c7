// K4 and the small helper kernels of the build: GetWordCodes per distinct word (Build.cs:175-247), host-supplied
// strings (Builder.AddWord / Builder.Add, header pages), page coordinates.  The tokenizer itself is in tokenize.cu.
#include "build_kernels.cuh"
#include "stem_en.cuh"
#include "stem_ru.cuh"

namespace dcd {

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

// Source of the words: slot list of the tokenizer's word table (slots != nullptr), or packed host-supplied strings
// (chars/offs; items with kinds[i] == 1 are raw codes and are skipped).  word_codes[i] = {n, entry ids of the codes}.
__global__ void __launch_bounds__(128) k_word_codes(BuildParams bp, WordTable wt, const u32* __restrict__ slots,
                                                     const u16* __restrict__ chars, const u32* __restrict__ offs,
                                                     const u8* __restrict__ kinds, u32 n_words, u32* __restrict__ word_codes) {
    const u32 idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_words) return;
    const StrTable& t = bp.tab;
    u32* out = word_codes + (u64)idx * 8;
    u16 w[MAX_KEY_UNITS + 2];
    int len;
    if (slots) {
        const ulonglong2 sl = wt.slots[slots[idx]];
        len = wordkey_unpack(WordKey{sl.x, sl.y}, wt.arena, w);
    } else {
        if (kinds && kinds[idx] == 1) { out[0] = 0; return; }
        len = (int)(offs[idx + 1] - offs[idx]);
        if (len > MAX_KEY_UNITS) len = MAX_KEY_UNITS;
        for (int i = 0; i < len; i++) w[i] = chars[offs[idx] + i];
    }
    if (len == 0) { out[0] = 0; return; }                      // Build.cs:178
    // the word itself is the first candidate code: find-or-insert it in the code table (stop words are already there)
    const u32 e = strtab_insert(t, w, len, hash_units(w, len), 0);
    if (e == ENT_NONE) { out[0] = 0; return; }
    const u32 flags = ld_volatile_u32(t.ent_flags + e);
    u32 n = 0;
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
void launch_word_codes(const BuildParams& bp, const WordTable& wt, const u32* slots, const u16* chars, const u32* offs,
                       const u8* kinds, u32 n_words, u32* word_codes, cudaStream_t s) {
    if (!n_words) return;
    k_word_codes<<<(n_words + 127) / 128, 128, 0, s>>>(bp, wt, slots, chars, offs, kinds, n_words, word_codes);
}

// after ranking: ordinal code ids -> slot_codes of the word table
__global__ void k_pack_slot_codes(const u32* __restrict__ slots, u32 n_words, const u32* __restrict__ word_codes,
                                  u64* __restrict__ slot_codes) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    const u32* r = word_codes + (u64)i * 8;
    const u32 n = r[0];
    u64 c;
    if (n <= 2) c = (u64)n | ((u64)(n > 0 ? r[1] : 0) << 4) | ((u64)(n > 1 ? r[2] : 0) << 28);
    else c = (u64)n | ((u64)i << 4);
    slot_codes[slots[i]] = c;
}
void launch_pack_slot_codes(const u32* slots, u32 n_words, const u32* word_codes, u64* slot_codes, cudaStream_t s) {
    if (!n_words) return;
    k_pack_slot_codes<<<(n_words + 255) / 256, 256, 0, s>>>(slots, n_words, word_codes, slot_codes);
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
// Strings of several packed tables in one launch (dictionary unification of the sharded build: the tables of all the other
// ranks at once -- millions of threads hide the probe latency that seven launches of ~0.6 M threads each showed).
__global__ void k_insert_strings_multi(StrTable t, const u8* __restrict__ chars_base, const u8* __restrict__ offs_base, InsertSources src,
                                       u32 flags) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= src.first[src.n]) return;
    int k = 0;
    while (i >= src.first[k + 1]) k++;
    const u16* chars = reinterpret_cast<const u16*>(chars_base + src.chars_off[k]);
    const u32* offs = reinterpret_cast<const u32*>(offs_base + src.offs_off[k]);
    const u32 j = i - src.first[k];
    const u16* s = chars + offs[j];
    const int len = (int)(offs[j + 1] - offs[j]);
    strtab_insert(t, s, len, hash_units(s, len), flags);
}
void launch_insert_strings_multi(const StrTable& tab, const void* chars_base, const void* offs_base, const InsertSources& src, u32 flags,
                                 cudaStream_t s) {
    const u32 n = src.first[src.n];
    if (!n) return;
    k_insert_strings_multi<<<(n + 127) / 128, 128, 0, s>>>(tab, (const u8*)chars_base, (const u8*)offs_base, src, flags);
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
    const u64 coord = coords[i] & COORD_MASK;
    if (kinds[i] == 1) {
        const u32 e = entries[i];
        if (e == ENT_NONE) return;
        keys[atomicAdd(nkeys, 1u)] = ((u64)t.ent_value[e] << COORD_BITS) | coord;
    } else {
        const u32* wc = word_codes + (u64)i * 8;
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
