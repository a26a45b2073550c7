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
void engine_search_batch(dcd_ctx_impl* c, const std::vector<u16s>& reqs, const std::vector<int>& has_opt, const std::vector<int>& dist,
                         dcd_hits* out);
}
