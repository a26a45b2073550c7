// K1-K3: single-pass tokenizer + distinct-word table + posting expansion.
//
// Replaces the per-token loop of IndexTask (Build.cs:477-536): ToLower (482), Regex \p{L}+|\p{N}+ (526), the
// 3..32 length filter (528), Builder.AddWord -> GetWordCodes (302-316, 175-247) and Builder.Add (324-341).
//
//   k_tokenize   text tile (8192 units, 32 per thread) -> SWAR lower-case + letter/digit BITMASKS -> token starts and
//                lengths by bit arithmetic -> exact 128-bit word key -> one 16-byte probe of the word table ->
//                token record (wordSlot << 40 | coord) at a fixed per-tile stride (coordinate order, no look-back).
//   K4           (build_kernels.cu) GetWordCodes once per DISTINCT word -> slot_codes[slot].
//   k_expand     token records -> posting keys (codeId << 40 | coord), dense and in coordinate order (decoupled
//                look-back), with the radix-sort digit histograms fused in.
#pragma once
#include "common.cuh"
#include "wordtable.cuh"

namespace dcd {

constexpr int TK_TILE = 8192;         // text units per tile
constexpr int TK_THREADS = 256;
constexpr int TK_CPT = TK_TILE / TK_THREADS;   // 32 units per thread == one mask word
constexpr int TK_LA = 64;             // look-ahead units (>= max word length + 1)
constexpr i64 PAGE_SKIP = (i64)0x8000000000000000ll;   // page_delta of pages the tokenizer must ignore (header pages)
constexpr i64 TILE_DELTA_MIXED = (i64)0x8000000000000001ll;   // tile_delta: pages of the tile differ -> per-token lookup

struct TextView {          // one batch of pages resident on the device
    const void* text;      // DCD_ENC_LATIN1 bytes or DCD_ENC_UTF16 units, 16-byte aligned
    int enc;
    u64 n_units;
    const u64* page_off;   // [npages + 1] ascending unit offsets
    const i64* page_delta; // [npages] coord = unit_offset + delta ; PAGE_SKIP for header pages
    const u32* tile_first_page;   // [ntiles] index of the page containing the first unit of the tile
    const u32* page_bits;  // [ntiles * 256 + 4] bit x: a page starts at unit x
    const i64* tile_delta; // [ntiles] common delta of every page that intersects the tile, or TILE_DELTA_MIXED
    u64 npages;
    u32 ntiles;
};

struct TokenizeParams {
    WordTable tab;
    const u8* cls;         // [65536] class bits
    const u16* lower;      // [65536] ToLower
    int min_len, max_len;
    u64* tok;              // [ntiles * tok_cap] token records of tile t at tok + t * tok_cap
    u32 tok_cap;           // >= TK_TILE / min_len + 1
    u32* tile_count;       // [ntiles]
    u64* token_total;      // tokens with min <= len <= max in non-skipped pages
    u32* error;            // [2] coordinate overflow
};

// slot_codes entry: bits [3:0] n codes; n <= 2: ids at [27:4] and [51:28]; n > 2: [35:4] index of an overflow
// record u32[8] = {n, id0..id5}
constexpr u64 SLOT_CODES_NONE = 0;

struct ExpandParams {
    const u64* tok;
    u32 tok_cap;
    const u32* tile_count;
    u32 ntiles;
    const u64* slot_codes;      // [table capacity]
    const u32* overflow;        // [n][8]
    u64* keys;                  // dense output
    u64 capacity;
    u64* lookback;              // [ntiles] zeroed
    u32* ticket;                // zeroed
    u64* hist;                  // [npasses][WIDE_HIST_STRIDE] (u64) digit histograms, zeroed
    int pass_shift[4];          // relative to the code id (bit 0 = lowest code-id bit)
    int pass_bits[4];
    int npasses;
    u64* total;                 // [0] postings emitted
    u32* error;                 // [0] capacity exceeded
};

void launch_tile_info(const u64* page_off, const i64* page_delta, u64 npages, u64 n_units, u32* tile_first_page, i64* tile_delta,
                      u32 ntiles, cudaStream_t s);
void launch_page_bits(const u64* page_off, u64 npages, u32* bits, cudaStream_t s);
size_t tokenize_smem_bytes(int enc, u32 tok_cap);
void launch_tokenize(const TextView& tv, const TokenizeParams& tp, int nsm, cudaStream_t s);
void launch_expand(const ExpandParams& ep, int nsm, cudaStream_t s);
// occupied slots of the table -> dense list (order unspecified); *out_n = count
void launch_list_words(const WordTable& tab, u32* out_slots, u32* out_n, cudaStream_t s);

}  // namespace dcd

// ---------------------------------------------------------------------------------------------------
// v2 tokenizer for batches whose pages share one coordinate delta (bulk pages): warp-autonomous, no block barriers.
// A warp classifies a 1024-unit window and owns the tokens that START in its first 960 units (the last 64 are the
// look-ahead, re-classified by the next tile).  Token records are 4 bytes: wordSlot:22 | start offset in the tile:10,
// at a fixed stride of T2_CAP records per tile; the consumer derives the coordinate as tile * 960 + offset + delta.
// ---------------------------------------------------------------------------------------------------
namespace dcd {
constexpr int T2_OWN = 960;
constexpr int T2_WIN = 1024;
constexpr int T2_CAP = 324;            // >= 960 / 3 + 1 records (needs min_word_length >= 3)
constexpr int T2_OFF_BITS = 10;
constexpr u32 T2_SLOT_NONE = 0x3FFFFFu;
constexpr int T2_WARPS = 8;

struct Tok2Params {
    const void* text;
    int enc;
    u64 n_units;
    const u32* page_bits;   // [n_units / 32 + 64]
    u32 ntiles;             // ceil(n_units / 960); a launch covers tiles [tile_begin, ntiles)
    u32 tile_begin;         // chunked launches: the text arrives from the host in pieces (engine.cu, K1)
    WordTable tab;          // capacity <= 2^22 slots
    const u8* cls;
    const u16* lower;
    int min_len, max_len;
    u32* rec;               // [ntiles * T2_CAP]
    u16* tile_count;        // [ntiles]
    u64* token_total;
};
void launch_tokenize2(const Tok2Params& tp, int nsm, cudaStream_t s);

constexpr int EX2_GROUP = 8;           // tiles per warp iteration of the expand kernel
struct Expand2Params {
    const u32* rec;
    const u16* tile_count;
    u32 ntiles;
    i64 delta;                  // coord = tile * 960 + offset + delta
    const u64* slot_codes;
    const u32* overflow;
    u64* keys;
    u64 capacity;
    u64* lookback;              // [ceil(ntiles / EX2_GROUP)] zeroed
    u32* ticket;
    u64* hist;                  // [npasses][WIDE_HIST_STRIDE]
    int pass_shift[4];
    int pass_bits[4];
    int npasses;
    u64* total;
    u32* error;
};
void launch_expand2(const Expand2Params& ep, int nsm, cudaStream_t s);
}  // namespace dcd
