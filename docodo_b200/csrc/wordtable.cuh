// Distinct-word table of the single-pass tokenizer (K2; replaces the per-token SortedDictionary<string,...> descent of
// Build.cs:328-332 and the per-token GetWordCodes call of Build.cs:305).
//
// One 16-byte slot per word; the slot IS the word: a 128-bit EXACT, invertible packing of the lower-cased token
//   tag 1  alpha : [a-z]{1,25}      5 bits per char (a=1 .. z=26), char j at bit 5j
//   tag 2  digit : [0-9]{1,30}      4 bits per char (0=1 .. 9=10), char j at bit 4j
//   tag 3  cyr   : [а-яё]{1,20}     6 bits per char (а=1 .. я=32, ё=33), char j at bit 6j
//   tag 4  other : lo = 64-bit hash, hi = tag | len << 32 | arena offset; the UTF-16 string is in `arena`
// (tag in bits 125..127).  A probe is one 128-bit load and one 128-bit compare -- no string compare, no second
// dependent access -- and an insert is one 128-bit CAS, so there is no "entry not ready yet" state to wait for.
// Tag-4 strings are written to the arena BEFORE the CAS that publishes their offset (a lost race only leaks the
// speculative arena bytes).  Equality of tag 1-3 keys is equality of the strings, so the table is exact, not
// probabilistic; tag 4 verifies the full string.
#pragma once
#include "common.cuh"
#include "strtable.cuh"

namespace dcd {

struct WordKey { u64 lo, hi; };
constexpr int WK_TAG_SHIFT = 61;
enum : u32 { WK_EMPTY = 0, WK_ALPHA = 1, WK_DIGIT = 2, WK_CYR = 3, WK_OTHER = 4, WK_SENTINEL = 7 /* reserved last slot */ };
constexpr u32 WORD_SLOT_NONE = 0xFFFFFFu;   // token record marker: no word (dropped token)

struct WordTable {
    ulonglong2* slots;   // [cap]
    u32 cap_mask;
    u16* arena;          // tag-4 strings
    u32 arena_cap;
    u32* counters;       // [0] arena units used, [1] overflow flag, [2] words inserted
};

DCD_HD u32 wordkey_tag(const WordKey& k) { return (u32)(k.hi >> WK_TAG_SHIFT); }

// slot hash: for tag 4 the arena offset (low 32 bits of hi) must not take part
DCD_HD u64 wordkey_slot_hash(const WordKey& k) {
    u64 hi = wordkey_tag(k) == WK_OTHER ? (k.hi & 0xFFFFFFFF00000000ull) : k.hi;
    u64 h = k.lo ^ (hi * 0x9E3779B97F4A7C15ull);
    h ^= h >> 32;
    h *= 0xD6E8FEB86659FD93ull;
    h ^= h >> 29;
    return h;
}

DCD_HD int cyr_code(u32 c) { return (c >= 0x0430 && c <= 0x044F) ? (int)(c - 0x0430 + 1) : (c == 0x0451 ? 33 : 0); }

// Generic (slow) packing from lower-cased UTF-16 units.  Returns the tag; for WK_OTHER only lo (hash) and the tag/len
// part of hi are filled -- the caller supplies the arena offset.
template <typename P>
DCD_HD u32 wordkey_pack(P s, int n, WordKey& k) {
    k.lo = k.hi = 0;
    if (n <= 0) return WK_EMPTY;
    bool alpha = n <= 25, digit = n <= 30, cyr = n <= 20;
    for (int i = 0; i < n; i++) {
        const u32 c = (u32)s[i];
        alpha = alpha && c >= 'a' && c <= 'z';
        digit = digit && c >= '0' && c <= '9';
        cyr = cyr && cyr_code(c) != 0;
    }
    if (alpha || digit || cyr) {
        const int bits = alpha ? 5 : (digit ? 4 : 6);
        for (int i = 0; i < n; i++) {
            const u32 c = (u32)s[i];
            const u64 v = alpha ? (u64)(c - 'a' + 1) : (digit ? (u64)(c - '0' + 1) : (u64)cyr_code(c));
            const int at = bits * i;
            if (at < 64) {
                k.lo |= v << at;
                if (at + bits > 64) k.hi |= v >> (64 - at);
            } else {
                k.hi |= v << (at - 64);
            }
        }
        const u32 tag = alpha ? WK_ALPHA : (digit ? WK_DIGIT : WK_CYR);
        k.hi |= (u64)tag << WK_TAG_SHIFT;
        return tag;
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
    const int bits = tag == WK_ALPHA ? 5 : (tag == WK_DIGIT ? 4 : 6);
    const u64 hi = k.hi & ((1ull << WK_TAG_SHIFT) - 1);
    int n = 0;
    for (int i = 0; i < 32; i++) {
        const int at = bits * i;
        if (at + bits > 64 + WK_TAG_SHIFT) break;
        u64 v;
        if (at >= 64) v = hi >> (at - 64);
        else {
            v = k.lo >> at;
            if (at + bits > 64) v |= hi << (64 - at);
        }
        v &= (1u << bits) - 1;
        if (!v) break;
        out[n++] = tag == WK_ALPHA ? (u16)('a' + v - 1) : (tag == WK_DIGIT ? (u16)('0' + v - 1) : (v == 33 ? (u16)0x0451 : (u16)(0x0430 + v - 1)));
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
    WordKey k{lo, hi};
    u32 i = (u32)wordkey_slot_hash(k) & t.cap_mask;
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
    WordKey k{hash, hi_tag};
    u32 i = (u32)wordkey_slot_hash(k) & t.cap_mask;
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
