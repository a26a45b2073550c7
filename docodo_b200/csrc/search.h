// Host side of Index.Search (Search.cs:437-617): request rewriting, planning into device programs, result shaping.
#pragma once
#include <string>
#include <vector>

#include "engine.h"

namespace dcd {

struct HitPage {
    i64 ordinal;
    float rank;
    u32 first;   // index into pos[]
    u32 npos;
};
struct HitDoc {
    i64 doc;
    float rank;
    int first_page;   // index into pages[] (pages of one document are contiguous)
    int npages;
};
struct HitWord {
    u16s word;
    int nfound;
};
struct QueryResult {
    int status = 0;   // DCD_OK, DCD_E_SYNTAX, DCD_E_STATE
    std::vector<HitPage> pages;
    std::vector<int32_t> pos;
    std::vector<HitDoc> docs;
    std::vector<HitWord> words;
};

}  // namespace dcd

struct dcd_hits {
    std::vector<dcd::QueryResult> q;
};

namespace dcd {
// Sharded serving: the found pages of a batch stay on the DEVICE as int32 pairs (page ordinal + page_base,
// positions in page); no per-request result objects are shaped on the host.  Filled by engine_search_batch, drained by
// engine_pages_fetch.
struct PagesSink {
    i64 page_base = 0;
    std::vector<int32_t> counts;                 // pages per request (after the MAX_FOUND_PAGES + 1 cut)
    std::vector<u64> req_src;                    // device address of the request's first pair
    std::vector<DevBuf<int32_t>> wave_pairs;     // per wave: the pairs of its programs, launch order
    u64 total = 0;
};
struct ShardComm;
// comm != nullptr (more than one rank): collective, exact serving from document-range shards -- every rank of the sharded
// build calls it with the same batch; bridging gap-runs of operator* are completed by a halo exchange (search.cu).
void engine_search_batch(dcd_ctx_impl* c, const std::vector<u16s>& reqs, const std::vector<int>& has_opt, const std::vector<int>& dist,
                         dcd_hits* out, PagesSink* sink = nullptr, ShardComm* comm = nullptr);
// request order, back to back, into dst (device pointer when on_device, else host); releases the sink's buffers
void engine_pages_fetch(dcd_ctx_impl* c, PagesSink& sink, int32_t* dst, u64 cap_pairs, bool on_device);
}
