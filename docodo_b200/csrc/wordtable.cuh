// Distinct-word table of the single-pass tokenizer (K2; replaces the per-token SortedDictionary<string,...> descent of
// Build.cs:328-332 and the per-token GetWordCodes call of Build.cs:305).
//
// One 16-byte slot per word; the slot IS the word: the lower-cased token as up to 15 raw bytes, zero padded
//   tag 1  bytes : 1..15 units, every unit U+0001..U+00FF (ASCII and Latin-1 words and numbers): byte j = unit j
//   tag 2  alpha : 16..24 units, every unit 'a'..'z': 5 bits per letter (a = 1), letter j at bit 5j (long words are rare in
//                  text but a dictionary-driven corpus can put one at a top Zipf rank)
//   tag 3  cyr   : 1..15 units, every unit below U+0080 or in the Cyrillic block U+0400..U+047F, at least one of the
//                  latter: byte j = unit < 0x80 ? unit : 0x80 | (unit & 0x7F)
//   tag 4  other : lo = 64-bit hash, hi = tag | len << 32 | arena offset; the UTF-16 string is in `arena`
// byte j sits at bit 8j of the 120-bit string lo | hi << 64, the tag in bits 125..127 of the slot.  The tokenizer keeps the
// lower-cased text of its tile as exactly these bytes in shared memory, so a key is two funnel shifts and a mask -- no
// per-letter packing.  A probe is one 128-bit load and one 128-bit compare (no string compare, no second dependent
// access); an insert is one 128-bit CAS, so there is no "entry not ready yet" state to wait for.  Tag-4 strings are
// written to the arena BEFORE the CAS that publishes their offset (a lost race only leaks the speculative arena bytes).
// Equality of tag 1 / 3 keys is equality of the strings (each string has exactly one representation), so the table is
// exact, not probabilistic; tag 4 verifies the full string.
#pragma once
#include "common.cuh"
#include "strtable.cuh"

namespace dcd {

struct WordKey { u64 lo, hi; };
constexpr int WK_TAG_SHIFT = 61;
constexpr int WK_MAX_EXACT = 15;   // units of an exact (tag 1 / 3) key
constexpr int WK_MAX_ALPHA = 24;   // letters of a tag-2 key
enum : u32 { WK_EMPTY = 0, WK_BYTES = 1, WK_ALPHA = 2, WK_CYR = 3, WK_OTHER = 4, WK_SENTINEL = 7 /* reserved last slot */ };
constexpr u32 WORD_SLOT_NONE = 0xFFFFFFu;    // "no word" (table full / empty string)

struct WordTable {
    ulonglong2* slots;   // [cap]
    u32 cap_mask;
    u16* arena;          // tag-4 strings
    u32 arena_cap;
    u32* counters;       // [0] arena units used, [1] overflow flag, [2] words inserted
};

DCD_HD u32 wordkey_tag(const WordKey& k) { return (u32)(k.hi >> WK_TAG_SHIFT); }

// slot hash (32-bit multiply-adds: the tokenizer computes it once per token).  For tag 4 the arena offset (low 32 bits of
// hi) must not take part.
DCD_HD u32 wordkey_slot_hash(u64 lo, u64 hi) {
    const u32 hi_lo = (u32)(hi >> WK_TAG_SHIFT) == WK_OTHER ? 0u : (u32)hi;
    u32 h = (u32)lo * 0x9E3779B1u + (u32)(lo >> 32) * 0x85EBCA77u + hi_lo * 0xC2B2AE3Du + (u32)(hi >> 32) * 0x27D4EB2Fu;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 13;
    return h;
}

DCD_HD bool is_cyr_block(u32 c) { return c >= 0x0400u && c < 0x0480u; }

// Generic (slow) packing from lower-cased UTF-16 units.  Returns the tag; for WK_OTHER only lo (hash) and the tag/len
// part of hi are filled -- the caller supplies the arena offset.
template <typename P>
DCD_HD u32 wordkey_pack(P s, int n, WordKey& k) {
    k.lo = k.hi = 0;
    if (n <= 0) return WK_EMPTY;
    bool bytes = n <= WK_MAX_EXACT, cyr = n <= WK_MAX_EXACT, anycyr = false;
    for (int i = 0; i < n; i++) {
        const u32 c = (u32)s[i];
        bytes = bytes && c != 0 && c < 0x100u;
        cyr = cyr && c != 0 && (c < 0x80u || is_cyr_block(c));
        anycyr = anycyr || is_cyr_block(c);
    }
    if (bytes || (cyr && anycyr)) {
        for (int i = 0; i < n; i++) {
            const u32 c = (u32)s[i];
            const u64 v = c < 0x100u ? (u64)c : (u64)(0x80u | (c & 0x7Fu));
            if (i < 8) k.lo |= v << (8 * i); else k.hi |= v << (8 * (i - 8));
        }
        const u32 tag = bytes ? WK_BYTES : WK_CYR;
        k.hi |= (u64)tag << WK_TAG_SHIFT;
        return tag;
    }
    if (n > WK_MAX_EXACT && n <= WK_MAX_ALPHA) {
        bool alpha = true;
        for (int i = 0; i < n; i++) alpha = alpha && (u32)s[i] >= 'a' && (u32)s[i] <= 'z';
        if (alpha) {
            for (int i = 0; i < n; i++) {
                const u64 v = (u64)((u32)s[i] - 'a' + 1);
                const int at = 5 * i;
                if (at < 64) {
                    k.lo |= v << at;
                    if (at + 5 > 64) k.hi |= v >> (64 - at);
                } else {
                    k.hi |= v << (at - 64);
                }
            }
            k.hi |= (u64)WK_ALPHA << WK_TAG_SHIFT;
            return WK_ALPHA;
        }
    }
    k.lo = hash_units(s, n);
    k.hi = ((u64)WK_OTHER << WK_TAG_SHIFT) | ((u64)n << 32);
    return WK_OTHER;
}

// Inverse: the UTF-16 string of a slot.  Returns the length.  `out` needs MAX_KEY_UNITS entries.
#if defined(__CUDACC__)
DCD_D int wordkey_unpack(const WordKey& k, const u16* arena, u16* out) {
    const u32 tag = wordkey_tag(k);
    if (tag == WK_OTHER) {
        const int n = (int)((k.hi >> 32) & 63);
        const u16* a = arena + (u32)k.hi;
        for (int i = 0; i < n; i++) out[i] = __ldcg(a + i);
        return n;
    }
    int n = 0;
    if (tag == WK_ALPHA) {
        const u64 hi = k.hi & ((1ull << WK_TAG_SHIFT) - 1);
        for (int i = 0; i < WK_MAX_ALPHA; i++) {
            const int at = 5 * i;
            u64 v;
            if (at >= 64) v = hi >> (at - 64);
            else {
                v = k.lo >> at;
                if (at + 5 > 64) v |= hi << (64 - at);
            }
            v &= 31;
            if (!v) break;
            out[n++] = (u16)('a' + v - 1);
        }
        return n;
    }
    for (int i = 0; i < WK_MAX_EXACT; i++) {
        const u32 b = (u32)((i < 8 ? k.lo >> (8 * i) : k.hi >> (8 * (i - 8))) & 0xFFu);
        if (!b) break;
        out[n++] = (tag == WK_CYR && b >= 0x80u) ? (u16)(0x0400u | (b & 0x7Fu)) : (u16)b;
    }
    return n;
}

DCD_D ulonglong2 ld_slot_ca(const ulonglong2* p) {
    ulonglong2 r;
    asm volatile("ld.global.ca.v2.u64 {%0,%1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
    return r;
}
DCD_D ulonglong2 ld_slot_cg(const ulonglong2* p) {
    ulonglong2 r;
    asm volatile("ld.global.cg.v2.u64 {%0,%1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p) : "memory");
    return r;
}
// 128-bit compare-and-swap of an empty slot (sm_90+: ATOMG.E.CAS.128).  Returns the previous content.
DCD_D ulonglong2 cas_slot_empty(ulonglong2* p, u64 nlo, u64 nhi) {
    ulonglong2 o;
    asm volatile(
        "{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%2, %3};\n\tmov.b128 n, {%4, %5};\n\t"
        "atom.global.relaxed.gpu.cas.b128 o, [%6], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
        : "=l"(o.x), "=l"(o.y)
        : "l"(0ull), "l"(0ull), "l"(nlo), "l"(nhi), "l"(p)
        : "memory");
    return o;
}

// Find-or-insert of an exact (tag 1-3) key.  Returns the slot index, or WORD_SLOT_NONE when the table is full.
// Slots only ever change from empty to their final value, so an L1-cached copy of a non-empty slot is always right;
// an apparently empty slot is re-read from L2 before the CAS.
DCD_D u32 wordtab_intern(const WordTable& t, u64 lo, u64 hi) {
    u32 i = wordkey_slot_hash(lo, hi) & t.cap_mask;
    for (u32 probe = 0; probe < 4096; probe++) {
        ulonglong2 s = ld_slot_ca(t.slots + i);
        if (s.x == lo && s.y == hi) return i;
        if (s.x == 0 || s.y == 0) {   // empty (or a torn view of a slot being written): look at L2
            s = ld_slot_cg(t.slots + i);
            if (s.x == 0 && s.y == 0) {
                s = cas_slot_empty(t.slots + i, lo, hi);
                if (s.x == 0 && s.y == 0) { atomicAdd(t.counters + 2, 1u); return i; }
            }
            if (s.x == lo && s.y == hi) return i;
        }
        i = (i + 1) & t.cap_mask;
    }
    atomicExch(t.counters + 1, 1u);
    return WORD_SLOT_NONE;
}

// Find-or-insert of a tag-4 word given its lower-cased units.
template <typename P>
DCD_D u32 wordtab_intern_other(const WordTable& t, P s, int n, u64 hash) {
    const u64 hi_tag = ((u64)WK_OTHER << WK_TAG_SHIFT) | ((u64)n << 32);
    u32 i = wordkey_slot_hash(hash, hi_tag) & t.cap_mask;
    u32 my_off = 0xFFFFFFFFu;
    for (u32 probe = 0; probe < 4096; probe++) {
        ulonglong2 sl = ld_slot_cg(t.slots + i);
        if (sl.x == 0 && sl.y == 0) {
            if (my_off == 0xFFFFFFFFu) {   // speculative arena copy, published by the CAS below
                const u32 pad = ((u32)n + 3u) & ~3u;
                my_off = atomicAdd(t.counters + 0, pad);
                if (my_off + pad > t.arena_cap) { atomicExch(t.counters + 1, 1u); return WORD_SLOT_NONE; }
                for (int j = 0; j < n; j++) t.arena[my_off + j] = (u16)s[j];
                __threadfence();
            }
            sl = cas_slot_empty(t.slots + i, hash, hi_tag | my_off);
            if (sl.x == 0 && sl.y == 0) { atomicAdd(t.counters + 2, 1u); return i; }
        }
        if (sl.x == hash && (sl.y & 0xFFFFFFFF00000000ull) == hi_tag) {
            __threadfence();
            const u16* a = t.arena + (u32)sl.y;
            bool eq = true;
            for (int j = 0; j < n && eq; j++) eq = __ldcg(a + j) == (u16)s[j];
            if (eq) return i;
        }
        i = (i + 1) & t.cap_mask;
    }
    atomicExch(t.counters + 1, 1u);
    return WORD_SLOT_NONE;
}

// Generic entry point for strings that do not come from the tokenizer (host-supplied AddWord arguments).
template <typename P>
DCD_D u32 wordtab_intern_units(const WordTable& t, P s, int n) {
    WordKey k;
    const u32 tag = wordkey_pack(s, n, k);
    if (tag == WK_EMPTY) return WORD_SLOT_NONE;
    if (tag != WK_OTHER) return wordtab_intern(t, k.lo, k.hi);
    return wordtab_intern_other(t, s, n, k.lo);
}
#endif

}  // namespace dcd
