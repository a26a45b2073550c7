// Build-side kernels K1-K4: tokenizer + distinct-word table + GetWordCodes + posting emit.
#pragma once
#include "common.cuh"
#include "strtable.cuh"

namespace dcd {

constexpr int TOK_TILE = 4096;       // UTF-16 units per tile
constexpr int TOK_THREADS = 256;
constexpr int TOK_PER_THREAD = TOK_TILE / TOK_THREADS;   // 16
constexpr int TOK_LA = MAX_KEY_UNITS + 1;                // look-ahead units needed to size a token
constexpr int TOK_MAX_TOKENS = TOK_TILE + 8;             // min_word_length may be 1
constexpr int TOK_PG_CACHE = 192;                        // page starts cached per tile
constexpr i64 PAGE_SKIP = (i64)0x8000000000000000ll;     // page_delta value of pages the tokenizer must ignore

struct TextView {          // one batch of pages resident on the device
    const void* text;      // DCD_ENC_LATIN1 bytes or DCD_ENC_UTF16 units, 16-byte aligned
    int enc;
    u64 n_units;
    const u64* page_off;   // [npages + 1] ascending unit offsets
    const i64* page_delta; // [npages] coord = unit_offset + delta ; PAGE_SKIP for header pages
    const u32* tile_first_page;   // [ntiles] index of the page containing the first unit of the tile
    u64 npages;
    u32 ntiles;
};

struct VocabDev {          // Vocab (Dict.cs:13-104) on the device
    StrTable tab;          // key -> group id in ent_value
    u16 begin, end;        // Range (Dict.cs:92-93)
    int stemmer;           // DCD_STEM_*
};

struct BuildParams {
    StrTable tab;          // main string table: words and codes
    const u8* cls;         // [65536] class bits
    const u16* lower;      // [65536] ToLower
    int min_len, max_len;
    int nvocs;
    VocabDev vocs[MAX_VOCS];
    int stemmers_mask;     // bit k: KnownStemmers[k] (digit, ru, en, de, fr) present in Index.Stemmers
};

struct EmitParams {
    u64* keys;             // [capacity] packed (codeId << 40 | coord), in coordinate order
    u64 capacity;
    const u32* word_codes; // [entries][8]: [0] = n, [1..] = ordinal code ids
    u64* lookback;         // [ntiles] decoupled look-back status
    u64* hist;             // [4][256] digit histograms of the code-id bits for the sort passes
    int pass_shift[4];
    int pass_bits[4];
    int npasses;
    u32* error;            // [0] key capacity exceeded, [1] unknown word, [2] coord/code overflow
    u64* total;            // [0] postings emitted, [1] tokens seen
};

void launch_tile_first_page(const u64* page_off, u64 npages, u32* tile_first_page, u32 ntiles, cudaStream_t s);
void launch_tokenize_insert(const TextView& tv, const BuildParams& bp, u32* ticket, u64* token_count, int nsm, cudaStream_t s);
void launch_tokenize_emit(const TextView& tv, const BuildParams& bp, const EmitParams& ep, u32* ticket, int nsm, cudaStream_t s);

// K4: GetWordCodes for every distinct word (entries [0, n_entries) with EF_WORD).  word_codes receives entry ids.
void launch_word_codes(const BuildParams& bp, u32 n_entries, u32* word_codes, cudaStream_t s);
// After ranking: replace entry ids in word_codes by ordinal code ids (ent_value).
void launch_codes_to_ids(const StrTable& tab, u32 n_entries, u32* word_codes, cudaStream_t s);
// Insert packed strings (host-supplied AddWord / Add arguments, stop words, vocabulary keys).
void launch_insert_strings(const StrTable& tab, const u16* chars, const u32* offs, u32 n, u32 flags, const u32* values,
                           u32* out_entries, cudaStream_t s);
// Gather entry ids having `flag` into out (order unspecified); count in *out_n.
void launch_gather_flagged(const StrTable& tab, u32 n_entries, u32 flag, u32* out, u32* out_n, cudaStream_t s);
// Postings of host-supplied words / raw codes: kind 0 = AddWord (expand through word_codes), 1 = Add (single code).
void launch_emit_extras(const StrTable& tab, const u32* word_codes, const u32* entries, const u64* coords, const u8* kinds,
                        u32 n, u64* keys, u32* nkeys, cudaStream_t s);

// Pages: adv[p] = page_off[p+1]-page_off[p] unless overridden; coordinates by scan.
void launch_page_adv(const u64* page_off, u64 npages, u64* adv, cudaStream_t s);
void launch_page_delta(const u64* page_off, const u64* adv, const u64* coord_end, const u8* kind, u64 npages, i64* delta,
                       cudaStream_t s);

}  // namespace dcd
