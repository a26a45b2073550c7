// Device-resident open-addressing string table (replaces the SortedDictionary<string,...> lookups of
// Build.cs:328-332, the Dictionary<string,int> of Dict.cs:13/98-103 and the stop-word HashSet of Index.cs:226).
//
// Layout (all in HBM, hot part lives in the 126 MB L2):
//   slot_hash[cap]   u64   0 = empty, else the 64-bit hash of the key (hash 0 is remapped to 1)
//   slot_entry[cap]  u32   dense entry id, 0xFFFFFFFF while the inserting thread is still writing the entry
//   ent_off/ent_len/ent_flags[max_entries]; arena[arena_cap] u16 -- the key characters (full-key verify)
// Dense entry ids make the later per-distinct-word work (GetWordCodes, ordinal ranking) a plain 1-D grid.
#pragma once
#include "common.cuh"

namespace dcd {

struct StrTable {
    u64* slot_hash;
    u32* slot_entry;
    u32* ent_off;
    u32* ent_len;     // length in UTF-16 units
    u32* ent_flags;
    u32* ent_value;   // per-entry payload (vocabulary: group id; main table: ordinal code id)
    u16* arena;
    u32* counters;    // [0] entries used, [1] arena units used, [2] overflow flag
    u32 cap_mask;
    u32 max_entries;
    u32 arena_cap;
};

constexpr u32 ENT_NOT_READY = 0xFFFFFFFFu;
constexpr u32 ENT_NONE = 0xFFFFFFFFu;

// FNV-1a over UTF-16 units + murmur3 finalizer.  Shared by host (vocabulary upload) and device.
DCD_HD u64 hash_init() { return 0xcbf29ce484222325ull; }
DCD_HD u64 hash_step(u64 h, u32 unit) { return (h ^ (u64)unit) * 0x100000001b3ull; }
DCD_HD u64 hash_final(u64 h) {
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 33;
    return h ? h : 1ull;
}
template <typename P>
DCD_HD u64 hash_units(P s, int n) {
    u64 h = hash_init();
    for (int i = 0; i < n; i++) h = hash_step(h, (u32)s[i]);
    return hash_final(h);
}

#if defined(__CUDACC__)
// CG = true: bypass the (non-coherent) L1 -- required while other CTAs are still inserting, because a stale
// 128-byte line holding a neighbouring entry may already sit in this SM's L1.
template <bool CG, typename P>
DCD_D bool strtab_equal(const StrTable& t, u32 e, P s, int n) {
    u32 len = CG ? __ldcg(t.ent_len + e) : t.ent_len[e];
    if (len != (u32)n) return false;
    // Entries start on 8-byte boundaries and are padded to four units (strtab_insert): whole 64-bit words are compared, all of
    // them loaded before the first comparison (a unit-by-unit loop with an early exit is one L2 round trip per unit).
    const u64* a = reinterpret_cast<const u64*>(t.arena + (CG ? __ldcg(t.ent_off + e) : t.ent_off[e]));
    u64 diff = 0;
    for (int w = 0; 4 * w < n; w++) {
        const u64 av = CG ? __ldcg(a + w) : a[w];
        u64 sv = 0;
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (4 * w + j < n) sv |= (u64)(u16)s[4 * w + j] << (16 * j);
        const u64 mask = 4 * w + 4 <= n ? ~0ull : (1ull << (16 * (n - 4 * w))) - 1;   // the padding of the last word is not written
        diff |= (av ^ sv) & mask;
    }
    return diff == 0;
}

// Lookup only.  Returns entry id or ENT_NONE.  Safe only when no insert is in flight.
template <typename P>
DCD_D u32 strtab_find(const StrTable& t, P s, int n, u64 h) {
    u32 i = (u32)h & t.cap_mask;
    for (u32 probe = 0; probe <= t.cap_mask; probe++) {
        u64 sh = t.slot_hash[i];
        if (sh == 0) return ENT_NONE;
        if (sh == h) {
            u32 e = t.slot_entry[i];
            if (e < t.max_entries && strtab_equal<false>(t, e, s, n)) return e;
        }
        i = (i + 1) & t.cap_mask;
    }
    return ENT_NONE;
}

// Find-or-insert; ORs `flags` into the entry.  Returns entry id, or ENT_NONE when the table is full
// (counters[2] is raised; the host turns that into DCD_E_CAPACITY).
template <typename P>
DCD_D u32 strtab_insert(const StrTable& t, P s, int n, u64 h, u32 flags) {
    u32 i = (u32)h & t.cap_mask;
    for (u32 probe = 0; probe <= t.cap_mask; probe++) {
        u64 sh = ld_volatile_u64(t.slot_hash + i);
        if (sh == 0) {
            u64 old = atomicCAS(t.slot_hash + i, 0ull, h);
            if (old == 0) {
                u32 e = atomicAdd(t.counters + 0, 1u);
                u32 pad = ((u32)n + 3u) & ~3u;
                u32 off = atomicAdd(t.counters + 1, pad);
                if (e >= t.max_entries || off + pad > t.arena_cap) {
                    atomicExch(t.counters + 2, 1u);
                    // leave the slot "not ready" forever would hang waiters: publish a sentinel entry 0-length
                    // is unsafe too, so publish ENT_NONE - 1 which readers treat as mismatch.
                    st_volatile_u32(t.slot_entry + i, ENT_NONE - 1);
                    return ENT_NONE;
                }
                u16* a = t.arena + off;
                for (int k = 0; k < n; k++) a[k] = (u16)s[k];
                t.ent_off[e] = off;
                t.ent_len[e] = (u32)n;
                t.ent_flags[e] = flags;
                t.ent_value[e] = ENT_NONE;
                st_release_u32(t.slot_entry + i, e);
                return e;
            }
            sh = old;
        }
        if (sh == h) {
            u32 e;
            while ((e = ld_acquire_u32(t.slot_entry + i)) == ENT_NOT_READY) __nanosleep(20);
            if (e != ENT_NONE - 1) {
                if (strtab_equal<true>(t, e, s, n)) {
                    if (flags && (ld_volatile_u32(t.ent_flags + e) & flags) != flags) atomicOr(t.ent_flags + e, flags);
                    return e;
                }
            }
        }
        i = (i + 1) & t.cap_mask;
    }
    atomicExch(t.counters + 2, 1u);
    return ENT_NONE;
}
#endif

}  // namespace dcd
