// ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle_core.h).  Flat C API for ctypes (tests, bench).
#include <chrono>
#include <cstring>
#include <thread>

#include "oracle_core.h"

using namespace orc;

namespace {
struct Handle {
    Index ix;
    std::vector<uint8_t> result;   // last serialized search result
    std::vector<uint16_t> scratch16;
    std::vector<uint64_t> scratch64;
};
u16s mk(const uint16_t* p, size_t n) { return u16s((const char16_t*)p, n); }
void put_i32(std::vector<uint8_t>& o, int32_t v) { uint8_t b[4]; memcpy(b, &v, 4); o.insert(o.end(), b, b + 4); }
void put_f32(std::vector<uint8_t>& o, float v) { uint8_t b[4]; memcpy(b, &v, 4); o.insert(o.end(), b, b + 4); }
void put_s16(std::vector<uint8_t>& o, const u16s& s) {
    put_i32(o, (int32_t)s.size());
    const uint8_t* p = (const uint8_t*)s.data();
    o.insert(o.end(), p, p + 2 * s.size());
}
u16s dec_ascii(int64_t v) { std::string t = std::to_string(v); return u16s(t.begin(), t.end()); }
}  // namespace

extern "C" {

void* orc_new() { return new Handle(); }
void orc_free(void* h) { delete (Handle*)h; }

int orc_add_voc_file(void* h, const uint8_t* bytes, size_t n, int stemmer) {
    auto v = std::make_shared<Vocab>();
    v->stemmer = stemmer;
    if (!v->load(bytes, n)) return -1;
    ((Handle*)h)->ix.vocs.push_back(v);
    return 0;
}
int orc_add_voc_pairs(void* h, const uint16_t* chars, const uint32_t* offs, const int32_t* vals, int n, uint16_t begin,
                      uint16_t end, int stemmer) {
    auto v = std::make_shared<Vocab>();
    v->stemmer = stemmer;
    for (int i = 0; i < n; i++) {
        v->map[mk(chars + offs[i], offs[i + 1] - offs[i])] = vals[i];
        v->ordered.emplace_back(mk(chars + offs[i], offs[i + 1] - offs[i]), vals[i]);
    }
    v->begin = begin;
    v->end = end;
    ((Handle*)h)->ix.vocs.push_back(v);
    return 0;
}
void orc_set_stopwords(void* h, const uint16_t* chars, const uint32_t* offs, int n) {
    auto& s = ((Handle*)h)->ix.stopWords;
    s.clear();
    for (int i = 0; i < n; i++) s.insert(mk(chars + offs[i], offs[i + 1] - offs[i]));
}
void orc_clear_stemmers(void* h) { ((Handle*)h)->ix.Stemmers.clear(); }
void orc_set_caps(void* h, int max_pages, int max_docs) {
    ((Handle*)h)->ix.MAX_FOUND_PAGES = max_pages;
    ((Handle*)h)->ix.MAX_FOUND_DOCS = max_docs;
}

void orc_add_doc(void* h, const uint16_t* src, int ns, const uint16_t* name, int nn) {
    ((Handle*)h)->ix.AddDoc(mk(src, ns), mk(name, nn));
}
void orc_index_page(void* h, const uint16_t* id, int nid, const uint16_t* text, size_t n) {
    ((Handle*)h)->ix.IndexPage(mk(id, nid), mk(text, n));
}
void orc_builder_add(void* h, const uint16_t* code, int n, uint64_t coord) { ((Handle*)h)->ix.Add(mk(code, n), coord); }
void orc_builder_add_word(void* h, const uint16_t* w, int n, uint64_t coord) { ((Handle*)h)->ix.AddWord(mk(w, n), coord); }
void orc_builder_end_page(void* h, const uint16_t* id, int n, int use_coord, uint64_t coord) {
    if (use_coord) ((Handle*)h)->ix.EndPage(mk(id, n), coord);
    else ((Handle*)h)->ix.EndPage(mk(id, n));
}
void orc_finish(void* h) { ((Handle*)h)->ix.Finish(); }
uint64_t orc_total_postings(void* h) { return ((Handle*)h)->ix.totalPostings; }
uint64_t orc_num_codes(void* h) { return ((Handle*)h)->ix.self.size(); }

size_t orc_index_bytes(void* h, const uint8_t** p) { *p = ((Handle*)h)->ix.index_bytes.data(); return ((Handle*)h)->ix.index_bytes.size(); }
size_t orc_list_bytes(void* h, const uint8_t** p) { *p = ((Handle*)h)->ix.list_bytes.data(); return ((Handle*)h)->ix.list_bytes.size(); }
int orc_load(void* h, const uint8_t* ib, size_t in, const uint8_t* lb, size_t ln) {
    return ((Handle*)h)->ix.Load(ib, in, lb, ln) ? 0 : -1;
}

// Bulk build used by the CPU baseline: pages are slices of one UTF-16 (enc=2) or single-byte
// ASCII/Latin-1 (enc=1) buffer, ids are "1".."k" within each document, documents are named
// "<source>:d<ordinal>".  nthreads > 1 mirrors MaxDegreeOfParallelism=N with DETERMINISTIC
// contiguous document ranges and a final MergeIndexes-style stitch (Index.cs:577-752) using exact
// text-length bases (canonical semantics; see SURVEY.md A.8 for why maxCoord shifts are not copied).
double orc_build_bulk(void* hv, const void* text, int enc, const uint64_t* page_offs, int64_t npages,
                      const uint64_t* doc_first_page, int64_t ndocs, int nthreads) {
    Handle* h = (Handle*)hv;
    auto t0 = std::chrono::steady_clock::now();
    if (nthreads < 1) nthreads = 1;
    if (nthreads > ndocs) nthreads = (int)std::max<int64_t>(1, ndocs);
    std::vector<std::unique_ptr<Index>> parts(nthreads);
    std::vector<uint64_t> part_len(nthreads, 0);
    auto work = [&](int t) {
        Index* ix = nthreads == 1 ? &h->ix : (parts[t].reset(new Index()), parts[t].get());
        if (nthreads != 1) {
            ix->vocs = h->ix.vocs;
            ix->stopWords = h->ix.stopWords;
            ix->Stemmers = h->ix.Stemmers;
        }
        int64_t d0 = ndocs * t / nthreads, d1 = ndocs * (t + 1) / nthreads;
        u16s page;
        for (int64_t d = d0; d < d1; d++) {
            ix->AddDoc(u"doc", u"d" + dec_ascii(d));
            for (uint64_t p = doc_first_page[d]; p < doc_first_page[d + 1]; p++) {
                size_t a = page_offs[p], b = page_offs[p + 1];
                if (enc == 2) page.assign((const char16_t*)text + a, b - a);
                else { page.resize(b - a); for (size_t k = 0; k < b - a; k++) page[k] = ((const uint8_t*)text)[a + k]; }
                ix->IndexPage(dec_ascii((int64_t)(p - doc_first_page[d] + 1)), page);
            }
        }
        part_len[t] = ix->coord;
    };
    if (nthreads == 1) {
        work(0);
        h->ix.Finish();
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
        // stitch: k-way merge by ordinal key, same-key lists concatenated in shard order with shift
        Index& out = h->ix;
        uint64_t base = 0;
        for (int t = 0; t < nthreads; t++) {
            Index& p = *parts[t];
            for (auto& kv : p.builder) {
                SeqBuilder& dst = out.builder[kv.first];
                for (uint64_t v : kv.second.seq.decode()) dst.add(v + base);   // Shift + AddRange, Index.cs:671-672
            }
            for (auto& pg : p.pages) out.pages.emplace_back(pg.first, pg.second + base);   // AddFromList(shift), Build.cs:53-72
            if (p.totalPostings) out.maxCoord = p.maxCoord + base;
            out.totalPostings += p.totalPostings;
            base += part_len[t];
        }
        out.coord = base;
        out.Finish();
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// ---- search ---------------------------------------------------------------------------------
size_t orc_search(void* hv, const uint16_t* req, int n, int has_opt, int dist, const uint8_t** out) {
    Handle* h = (Handle*)hv;
    SearchResult r = h->ix.Search(mk(req, n), has_opt != 0, dist);
    auto& o = h->result;
    o.clear();
    put_i32(o, r.success ? 1 : 0);
    put_i32(o, (int32_t)r.error.size());
    o.insert(o.end(), r.error.begin(), r.error.end());
    put_i32(o, (int32_t)r.foundPages.size());
    for (auto& p : r.foundPages) {
        put_i32(o, p.page_index);
        put_s16(o, p.id);
        put_f32(o, p.rank());
        put_i32(o, (int32_t)p.pos.size());
        for (int32_t v : p.pos) put_i32(o, v);
    }
    put_i32(o, (int32_t)r.foundDocs.size());
    for (auto& d : r.foundDocs) {
        put_i32(o, d.doc);
        put_s16(o, d.name);
        put_f32(o, d.rank);
        put_i32(o, (int32_t)d.pages.size());
        for (int v : d.pages) put_i32(o, v);
    }
    put_i32(o, (int32_t)r.words.size());
    for (auto& w : r.words) { put_s16(o, w.word); put_i32(o, w.nFound); }
    *out = o.data();
    return o.size();
}

// ---- small helpers for unit tests -----------------------------------------------------------
size_t orc_encode(void* hv, const uint64_t* v, size_t n, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    SeqBuilder b;
    for (size_t i = 0; i < n; i++) b.add(v[i]);
    h->scratch16 = b.seq.self;
    *out = h->scratch16.data();
    return h->scratch16.size();
}
size_t orc_decode(void* hv, const uint16_t* w, size_t n, const uint64_t** out) {
    Handle* h = (Handle*)hv;
    Seq s;
    s.self.assign(w, w + n);
    h->scratch64 = s.decode();
    *out = h->scratch64.data();
    return h->scratch64.size();
}
size_t orc_shift(void* hv, const uint16_t* w, size_t n, uint64_t shift, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    Seq s;
    s.self.assign(w, w + n);
    s.shift(shift);
    h->scratch16 = s.self;
    *out = h->scratch16.data();
    return h->scratch16.size();
}
// op: 0 = '*', 1 = '+'.  Returns the encoded result and its R.
size_t orc_op(void* hv, int op, const uint16_t* a, size_t na, int Ra, const uint16_t* b, size_t nb, int Rb, int* Rout,
              const uint16_t** out) {
    Handle* h = (Handle*)hv;
    Seq A, B;
    A.self.assign(a, a + na); A.R = Ra;
    B.self.assign(b, b + nb); B.R = Rb;
    Seq r = op == 0 ? op_mul(A, B) : op_add(A, B);
    *Rout = r.R;
    h->scratch16 = r.self;
    *out = h->scratch16.data();
    return h->scratch16.size();
}
size_t orc_stem(void* hv, int kind, const uint16_t* w, size_t n, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    tables_init();
    u16s r = kind == 1 ? stem_english(mk(w, n)) : kind == 2 ? stem_russian(mk(w, n)) : mk(w, n);
    h->scratch16.assign((const uint16_t*)r.data(), (const uint16_t*)r.data() + r.size());
    *out = h->scratch16.data();
    return r.size();
}
// codes joined by U+0000
size_t orc_word_codes(void* hv, const uint16_t* w, size_t n, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    auto codes = h->ix.GetWordCodes(mk(w, n));
    h->scratch16.clear();
    for (size_t i = 0; i < codes.size(); i++) {
        if (i) h->scratch16.push_back(0);
        h->scratch16.insert(h->scratch16.end(), (const uint16_t*)codes[i].data(), (const uint16_t*)codes[i].data() + codes[i].size());
    }
    *out = h->scratch16.data();
    return h->scratch16.size();
}
size_t orc_lower(void* hv, const uint16_t* w, size_t n, int upper, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    tables_init();
    u16s r = upper ? to_upper(mk(w, n)) : to_lower(mk(w, n));
    h->scratch16.assign((const uint16_t*)r.data(), (const uint16_t*)r.data() + r.size());
    *out = h->scratch16.data();
    return r.size();
}
int orc_char_class(uint16_t c) { tables_init(); return g_cls[c]; }

// ---- dictionary-side queries: strings joined by U+0000 ------------------------------------------
static size_t join16(Handle* h, const std::vector<u16s>& v, const uint16_t** out) {
    h->scratch16.clear();
    for (size_t i = 0; i < v.size(); i++) {
        if (i) h->scratch16.push_back(0);
        h->scratch16.insert(h->scratch16.end(), (const uint16_t*)v[i].data(), (const uint16_t*)v[i].data() + v[i].size());
    }
    *out = h->scratch16.data();
    return h->scratch16.size();
}
size_t orc_like_words(void* hv, const uint16_t* w, size_t n, int* count, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    auto v = h->ix.GetLikeWords(mk(w, n));
    *count = (int)v.size();
    return join16(h, v, out);
}
size_t orc_close_words(void* hv, const uint16_t* w, size_t n, int* count, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    auto v = h->ix.GetCloseWords(mk(w, n));
    *count = (int)v.size();
    return join16(h, v, out);
}
size_t orc_suggestions(void* hv, const uint16_t* req, size_t n, int nmax, int* count, const uint16_t** out) {
    Handle* h = (Handle*)hv;
    auto v = h->ix.GetSuggessions(mk(req, n), nmax);
    *count = (int)v.size();
    return join16(h, v, out);
}
size_t orc_histogram(void* hv, int top, int* count, const uint16_t** out, const uint64_t** counts) {
    Handle* h = (Handle*)hv;
    auto v = h->ix.CalcHistogram((size_t)top);
    std::vector<u16s> keys;
    h->scratch64.clear();
    for (auto& kv : v) { keys.push_back(kv.first); h->scratch64.push_back((uint64_t)kv.second); }
    *count = (int)v.size();
    *counts = h->scratch64.data();
    return join16(h, keys, out);
}
// ---- VocBuilder (Dict.cs:109-296) ------------------------------------------------------------------
void* orc_vb_new(int stemmer) { tables_init(); auto* b = new VocBuilder(); b->stemmer = stemmer; return b; }
void orc_vb_free(void* b) { delete (VocBuilder*)b; }
void orc_vb_add_group(void* b, const uint16_t* chars, const uint32_t* offs, int n) {
    std::vector<u16s> g;
    for (int i = 0; i < n; i++) g.push_back(mk(chars + offs[i], offs[i + 1] - offs[i]));
    ((VocBuilder*)b)->AddWordsGroup(g);
}
size_t orc_vb_build(void* b, uint8_t* out, size_t cap) {
    auto v = ((VocBuilder*)b)->build();
    if (out && v.size() <= cap) memcpy(out, v.data(), v.size());
    return v.size();
}

}  // extern "C"
