// Synthetic Zipf corpus generator (BASELINE.json configs 3-5).  Benchmark utility: produces the
// device-resident input of the build path; it is not part of the Docodo API and is never timed.
#pragma once
#include "common.cuh"

namespace dcd {
struct SynthParams {
    const u8* words;        // packed ASCII words in rank order
    const u32* word_offs;   // [nwords + 1]
    const double* cdf;      // [nwords] cumulative Zipf probabilities
    u32 nwords;
    u64 seed;
    u64 first_doc;          // global ordinal of the first generated document (document-range sharding)
    u64 ndocs;
    u32 pages_per_doc;
    u32 min_page_chars;
};
// pass 1: page_len[doc * pages_per_doc + p]; pass 2: text bytes at page_offs
void launch_synth_lengths(const SynthParams& sp, u64* page_len, cudaStream_t s);
void launch_synth_write(const SynthParams& sp, const u64* page_offs, u8* text, cudaStream_t s);
void launch_synth_doc_pages(u64 ndocs, u32 pages_per_doc, u64* doc_first_page, cudaStream_t s);
}  // namespace dcd
