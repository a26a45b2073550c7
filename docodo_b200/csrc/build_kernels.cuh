// Build-side kernels: K4 GetWordCodes per distinct word + helpers (the tokenizer K1-K3 is in tokenize.cuh).
#pragma once
#include "common.cuh"
#include "strtable.cuh"
#include "tokenize.cuh"

namespace dcd {

struct VocabDev {          // Vocab (Dict.cs:13-104) on the device
    StrTable tab;          // key -> group id in ent_value
    u16 begin, end;        // Range (Dict.cs:92-93)
    int stemmer;           // DCD_STEM_*
};

struct BuildParams {
    StrTable tab;          // code table: stop words and every code string ("word", "#HEX", "$stem", "&field")
    int nvocs;
    VocabDev vocs[MAX_VOCS];
    int stemmers_mask;     // bit k: KnownStemmers[k] (digit, ru, en, de, fr) present in Index.Stemmers
};

// K4: GetWordCodes for n_words distinct words taken from the word table (slots) or from packed strings (chars/offs,
// kinds[i] == 1 skipped).  word_codes[i] = {n, entry ids in bp.tab}.
void launch_word_codes(const BuildParams& bp, const WordTable& wt, const u32* slots, const u16* chars, const u32* offs,
                       const u8* kinds, u32 n_words, u32* word_codes, cudaStream_t s);
// After ranking: replace entry ids in word_codes by ordinal code ids (ent_value).
void launch_codes_to_ids(const StrTable& tab, u32 n_entries, u32* word_codes, cudaStream_t s);
// ... and publish them per word-table slot for the expand kernel.
void launch_pack_slot_codes(const u32* slots, u32 n_words, const u32* word_codes, u64* slot_codes, cudaStream_t s);
// Insert packed strings (host-supplied AddWord / Add arguments, stop words, vocabulary keys).
// up to 16 packed string tables inside two buffers: table k holds first[k + 1] - first[k] strings, its chars start at byte
// chars_off[k] of chars_base and its (count + 1) u32 offsets at byte offs_off[k] of offs_base
struct InsertSources { u32 first[17]; u64 chars_off[16], offs_off[16]; int n; };
void launch_insert_strings_multi(const StrTable& tab, const void* chars_base, const void* offs_base, const InsertSources& src, u32 flags,
                                 cudaStream_t s);
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
