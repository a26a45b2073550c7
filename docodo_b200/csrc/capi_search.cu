// extern "C" entry points for queries and the stand-alone IndexSequence algebra.
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#include "encode.cuh"
#include "engine.h"
#include "query_kernels.cuh"
#include "search.h"
#include "shard.h"

using namespace dcd;


template <typename F>
static int32_t guard2(dcd_ctx_impl* c, F&& f) {
    try {
        DCD_CUDA(cudaSetDevice(c->device));
        f();
        return DCD_OK;
    } catch (const StatusError& e) {
        c->err = e.what();
        return e.code;
    } catch (const CudaError& e) {
        cudaGetLastError();
        c->err = e.what();
        return DCD_E_CUDA;
    } catch (const std::exception& e) {
        c->err = e.what();
        return DCD_E_INVALID;
    }
}

static int32_t search_batch_impl(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* offs, int32_t nreq,
                                 const int32_t* has_opt, const int32_t* dist, dcd_hits** out);
extern "C" int32_t dcd_search_batch(dcd_ctx* c, const uint16_t* reqs, const uint32_t* offs, int32_t nreq, const int32_t* has_opt,
                                    const int32_t* dist, dcd_hits** out) {
    return search_batch_impl(c, nullptr, reqs, offs, nreq, has_opt, dist, out);
}
extern "C" int32_t dcd_search_batch_sharded(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* offs, int32_t nreq,
                                            const int32_t* has_opt, const int32_t* dist, dcd_hits** out) {
    if (!comm || !comm->impl) return DCD_E_INVALID;
    return search_batch_impl(c, comm, reqs, offs, nreq, has_opt, dist, out);
}
static int32_t search_batch_impl(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* offs, int32_t nreq,
                                 const int32_t* has_opt, const int32_t* dist, dcd_hits** out) {
    if (!c || !out || nreq < 0 || (nreq && (!reqs || !offs))) return DCD_E_INVALID;
    *out = nullptr;
    dcd_hits* h = new dcd_hits();
    int32_t st = guard2(c, [&] {
        std::vector<u16s> r((size_t)nreq);
        std::vector<int> ho((size_t)nreq), di((size_t)nreq);
        for (int i = 0; i < nreq; i++) {
            r[i] = u16s((const char16_t*)reqs + offs[i], offs[i + 1] - offs[i]);
            ho[i] = has_opt ? has_opt[i] : 0;
            di[i] = dist ? dist[i] : 0;
        }
        try {
            engine_search_batch(c, r, ho, di, h, nullptr, comm ? comm->impl : nullptr);
        } catch (...) {
            if (comm) comm->impl->abort();   // release the peers blocked in a collective
            throw;
        }
    });
    if (st != DCD_OK) { delete h; return st; }
    *out = h;
    return DCD_OK;
}
// Sharded serving: evaluate the batch on this shard and keep the found pages on the device (no result objects are shaped).
static int32_t search_pages_impl(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* offs, int32_t nreq,
                                 const int32_t* has_opt, const int32_t* dist, int64_t page_base, int32_t* page_counts, uint64_t* total);
extern "C" int32_t dcd_search_batch_pages(dcd_ctx* c, const uint16_t* reqs, const uint32_t* offs, int32_t nreq, const int32_t* has_opt,
                                          const int32_t* dist, int64_t page_base, int32_t* page_counts, uint64_t* total) {
    return search_pages_impl(c, nullptr, reqs, offs, nreq, has_opt, dist, page_base, page_counts, total);
}
extern "C" int32_t dcd_search_batch_pages_sharded(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* offs, int32_t nreq,
                                                  const int32_t* has_opt, const int32_t* dist, int64_t page_base, int32_t* page_counts,
                                                  uint64_t* total) {
    if (!comm || !comm->impl) return DCD_E_INVALID;
    return search_pages_impl(c, comm, reqs, offs, nreq, has_opt, dist, page_base, page_counts, total);
}
static int32_t search_pages_impl(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* offs, int32_t nreq,
                                 const int32_t* has_opt, const int32_t* dist, int64_t page_base, int32_t* page_counts, uint64_t* total) {
    if (!c || nreq < 0 || (nreq && (!reqs || !offs))) return DCD_E_INVALID;
    return guard2(c, [&] {
        std::vector<u16s> r((size_t)nreq);
        std::vector<int> ho((size_t)nreq), di((size_t)nreq);
        for (int i = 0; i < nreq; i++) {
            r[i] = u16s((const char16_t*)reqs + offs[i], offs[i + 1] - offs[i]);
            ho[i] = has_opt ? has_opt[i] : 0;
            di[i] = dist ? dist[i] : 0;
        }
        c->pages_sink.reset();
        auto sink = std::make_shared<PagesSink>();
        sink->page_base = page_base;
        dcd_hits scratch;
        try {
            engine_search_batch(c, r, ho, di, &scratch, sink.get(), comm ? comm->impl : nullptr);
        } catch (...) {
            if (comm) comm->impl->abort();
            throw;
        }
        if (sink->counts.size() != (size_t)nreq) sink->counts.assign((size_t)nreq, 0);   // index not built: no pages
        if (page_counts) memcpy(page_counts, sink->counts.data(), (size_t)nreq * 4);
        if (total) *total = sink->total;
        c->pages_sink = sink;
    });
}
extern "C" int32_t dcd_search_pages_fetch(dcd_ctx* c, int32_t* pairs, uint64_t cap_pairs, int32_t on_device) {
    if (!c) return DCD_E_INVALID;
    return guard2(c, [&] {
        if (!c->pages_sink) throw StatusError(DCD_E_STATE, "no pending pages: call dcd_search_batch_pages first");
        auto sink = std::static_pointer_cast<PagesSink>(c->pages_sink);
        if (sink->total && !pairs) throw StatusError(DCD_E_INVALID, "pairs is NULL");
        engine_pages_fetch(c, *sink, pairs, cap_pairs, on_device != 0);
        c->pages_sink.reset();
    });
}
extern "C" void dcd_hits_free(dcd_hits* h) { delete h; }
#define DCD_Q(h, q) if (!(h) || (q) < 0 || (size_t)(q) >= (h)->q.size()) return DCD_E_INVALID; const QueryResult& r = (h)->q[q]
extern "C" int32_t dcd_hits_status(const dcd_hits* h, int32_t q) { DCD_Q(h, q); return r.status; }
extern "C" int32_t dcd_hits_num_pages(const dcd_hits* h, int32_t q) { DCD_Q(h, q); return (int32_t)r.pages.size(); }
extern "C" int32_t dcd_hits_page(const dcd_hits* h, int32_t q, int32_t k, int64_t* ordinal, float* rank, const int32_t** pos, int32_t* npos) {
    DCD_Q(h, q);
    if (k < 0 || (size_t)k >= r.pages.size()) return DCD_E_INVALID;
    const HitPage& p = r.pages[k];
    if (ordinal) *ordinal = p.ordinal;
    if (rank) *rank = p.rank;
    if (pos) *pos = r.pos.data() + p.first;
    if (npos) *npos = (int32_t)p.npos;
    return DCD_OK;
}
// Flat export of every request's found pages (sharded serving: the per-shard lists are gathered and concatenated in rank order)
extern "C" int32_t dcd_hits_export_pages(const dcd_hits* h, int32_t* page_counts, int64_t* ordinals, float* ranks, int32_t* npos,
                                         uint64_t cap, uint64_t* total) {
    if (!h) return DCD_E_INVALID;
    uint64_t n = 0;
    for (size_t q = 0; q < h->q.size(); q++) {
        const QueryResult& r = h->q[q];
        if (page_counts) page_counts[q] = (int32_t)r.pages.size();
        if (ordinals || ranks || npos) {
            if (n + r.pages.size() > cap) return DCD_E_INVALID;
            for (size_t k = 0; k < r.pages.size(); k++) {
                if (ordinals) ordinals[n + k] = r.pages[k].ordinal;
                if (ranks) ranks[n + k] = r.pages[k].rank;
                if (npos) npos[n + k] = (int32_t)r.pages[k].npos;
            }
        }
        n += r.pages.size();
    }
    if (total) *total = n;
    return DCD_OK;
}
// Same export as (global page ordinal, positions in page) int32 pairs, the payload the sharded gather moves: one pass, the
// requests spread over host threads.  pairs may be page-locked memory (it then goes to the device without staging).
extern "C" int32_t dcd_hits_export_pairs32(const dcd_hits* h, int64_t page_base, int32_t* page_counts, int32_t* pairs, uint64_t cap,
                                           uint64_t* total) {
    if (!h) return DCD_E_INVALID;
    const size_t nq = h->q.size();
    std::vector<uint64_t> start(nq + 1, 0);
    for (size_t q = 0; q < nq; q++) {
        start[q + 1] = start[q] + h->q[q].pages.size();
        if (page_counts) page_counts[q] = (int32_t)h->q[q].pages.size();
    }
    if (total) *total = start[nq];
    if (!pairs) return DCD_OK;
    if (start[nq] > cap) return DCD_E_INVALID;
    const unsigned nt = (unsigned)std::max<size_t>(1, std::min<size_t>({(size_t)8, (size_t)std::thread::hardware_concurrency(), start[nq] / 65536 + 1}));
    auto work = [&](unsigned t) {
        // thread t takes the requests whose first page falls into its equal share of the pages
        const uint64_t lo = start[nq] * t / nt, hi = start[nq] * (t + 1) / nt;
        size_t q = std::lower_bound(start.begin(), start.end() - 1, lo) - start.begin();
        for (; q < nq && start[q] < hi; q++) {
            if (start[q] < lo) continue;
            const QueryResult& r = h->q[q];
            int32_t* o = pairs + 2 * start[q];
            for (size_t k = 0; k < r.pages.size(); k++) {
                o[2 * k] = (int32_t)(r.pages[k].ordinal + page_base);
                o[2 * k + 1] = (int32_t)r.pages[k].npos;
            }
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; t++) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
    return DCD_OK;
}
extern "C" int32_t dcd_hits_export_positions(const dcd_hits* h, int32_t q, int32_t* pos, uint64_t cap, uint64_t* total) {
    DCD_Q(h, q);
    uint64_t n = 0;
    for (const HitPage& p : r.pages) {
        if (pos) {
            if (n + p.npos > cap) return DCD_E_INVALID;
            memcpy(pos + n, r.pos.data() + p.first, (size_t)p.npos * 4);
        }
        n += p.npos;
    }
    if (total) *total = n;
    return DCD_OK;
}
extern "C" int32_t dcd_hits_num_docs(const dcd_hits* h, int32_t q) { DCD_Q(h, q); return (int32_t)r.docs.size(); }
extern "C" int32_t dcd_hits_doc(const dcd_hits* h, int32_t q, int32_t k, int64_t* doc, float* rank, int32_t* first_page, int32_t* npages) {
    DCD_Q(h, q);
    if (k < 0 || (size_t)k >= r.docs.size()) return DCD_E_INVALID;
    const HitDoc& d = r.docs[k];
    if (doc) *doc = d.doc;
    if (rank) *rank = d.rank;
    if (first_page) *first_page = d.first_page;
    if (npages) *npages = d.npages;
    return DCD_OK;
}
extern "C" int32_t dcd_hits_num_words(const dcd_hits* h, int32_t q) { DCD_Q(h, q); return (int32_t)r.words.size(); }
extern "C" int32_t dcd_hits_word(const dcd_hits* h, int32_t q, int32_t k, const uint16_t** word, int32_t* n, int32_t* nfound) {
    DCD_Q(h, q);
    if (k < 0 || (size_t)k >= r.words.size()) return DCD_E_INVALID;
    if (word) *word = (const uint16_t*)r.words[k].word.data();
    if (n) *n = (int32_t)r.words[k].word.size();
    if (nfound) *nfound = r.words[k].nfound;
    return DCD_OK;
}

// ---- dictionary-side queries over the resident key table (dict_kernels.cu) ---------------------------------------------
static int32_t join_out(dcd_ctx* c, const std::vector<u16s>& v, uint16_t* out, int32_t cap, int32_t* out_units, int32_t* count) {
    size_t total = 0;
    for (size_t i = 0; i < v.size(); i++) total += v[i].size() + (i ? 1 : 0);
    if (out_units) *out_units = (int32_t)total;
    if (count) *count = (int32_t)v.size();
    if (!out) return DCD_OK;
    if ((size_t)cap < total) { c->err = "output buffer too small"; return DCD_E_INVALID; }
    size_t at = 0;
    for (size_t i = 0; i < v.size(); i++) {
        if (i) out[at++] = 0;
        memcpy(out + at, v[i].data(), v[i].size() * 2);
        at += v[i].size();
    }
    return DCD_OK;
}
extern "C" int32_t dcd_like_words(dcd_ctx* c, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_units, int32_t* count) {
    if (!c || n < 0 || (n && !word)) return DCD_E_INVALID;
    std::vector<u16s> v;
    int32_t st = guard2(c, [&] { v = engine_like_words(c, u16s((const char16_t*)word, n)); });
    return st != DCD_OK ? st : join_out(c, v, out, cap, out_units, count);
}
extern "C" int32_t dcd_close_words(dcd_ctx* c, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_units, int32_t* count) {
    if (!c || n < 0 || (n && !word)) return DCD_E_INVALID;
    std::vector<u16s> v;
    int32_t st = guard2(c, [&] { v = engine_close_words(c, u16s((const char16_t*)word, n)); });
    return st != DCD_OK ? st : join_out(c, v, out, cap, out_units, count);
}
extern "C" int32_t dcd_suggestions(dcd_ctx* c, const uint16_t* req, int32_t n, int32_t max_words, uint16_t* out, int32_t cap,
                                   int32_t* out_units, int32_t* count) {
    if (!c || n < 0 || (n && !req)) return DCD_E_INVALID;
    std::vector<u16s> v;
    int32_t st = guard2(c, [&] { v = engine_suggestions(c, u16s((const char16_t*)req, n), max_words); });
    return st != DCD_OK ? st : join_out(c, v, out, cap, out_units, count);
}
extern "C" int32_t dcd_histogram(dcd_ctx* c, int32_t top, uint16_t* out, int32_t cap, int32_t* out_units, uint64_t* counts, int32_t* count) {
    if (!c) return DCD_E_INVALID;
    std::vector<std::pair<u16s, u64>> h;
    int32_t st = guard2(c, [&] { h = engine_histogram(c, top); });
    if (st != DCD_OK) return st;
    std::vector<u16s> v;
    for (size_t i = 0; i < h.size(); i++) { v.push_back(h[i].first); if (counts && out) counts[i] = h[i].second; }
    return join_out(c, v, out, cap, out_units, count);
}

// ---- IndexSequence algebra on caller lists ---------------------------------------------------------
extern "C" int32_t dcd_seq_encode(dcd_ctx* c, const uint64_t* values, uint64_t n, uint16_t* out, uint64_t out_cap, uint64_t* out_n) {
    if (!c || (!values && n) || !out_n) return DCD_E_INVALID;
    return guard2(c, [&] {
        // Builder.Add over one list == the K6/K7 encoder with a single code id 0
        cudaStream_t s = c->stream;
        DevBuf<u64> keys(c->pool, n + 1), lb(c->pool, encode_scratch_tiles(n) + 2), code_off(c->pool, 2), tot(c->pool, 2);
        DevBuf<u16> arena(c->pool, n * 5 + 8);
        DevBuf<u32> ticket(c->pool, 16), err(c->pool, 16);
        std::vector<u64> k(n);
        for (u64 i = 0; i < n; i++) {
            if (values[i] > COORD_MASK) throw StatusError(DCD_E_UNSUPPORTED, "dcd_seq_encode handles coordinates below 2^40");
            k[i] = values[i];
        }
        *out_n = 0;
        if (!n) return;
        DCD_CUDA(cudaMemcpyAsync(keys.p, k.data(), n * 8, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
        DCD_CUDA(cudaMemsetAsync(err.p, 0, 64, s));
        launch_encode(keys.p, n, 1, arena.p, n * 5, code_off.p, nullptr, nullptr, lb.p, ticket.p, tot.p, nullptr, err.p, s);
        u64 total = 0;
        DCD_CUDA(cudaMemcpyAsync(&total, tot.p, 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        *out_n = total;
        if (total > out_cap) throw StatusError(DCD_E_INVALID, "output buffer too small");
        DCD_CUDA(cudaMemcpy(out, arena.p, total * 2, cudaMemcpyDeviceToHost));
    });
}
extern "C" int32_t dcd_seq_decode(dcd_ctx* c, const uint16_t* words, uint64_t n, uint64_t* out, uint64_t out_cap, uint64_t* out_n) {
    if (!c || (!words && n) || !out_n) return DCD_E_INVALID;
    return guard2(c, [&] {
        cudaStream_t s = c->stream;
        *out_n = 0;
        if (!n) return;
        DevBuf<u16> w(c->pool, n);
        DevBuf<u64> v(c->pool, n);
        DevBuf<u32> cnt(c->pool, 4);
        DCD_CUDA(cudaMemcpyAsync(w.p, words, n * 2, cudaMemcpyHostToDevice, s));
        launch_seq_decode(w.p, n, v.p, cnt.p, s);
        u32 m = 0;
        DCD_CUDA(cudaMemcpyAsync(&m, cnt.p, 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        *out_n = m;
        if (m > out_cap) throw StatusError(DCD_E_INVALID, "output buffer too small");
        if (m) DCD_CUDA(cudaMemcpy(out, v.p, (size_t)m * 8, cudaMemcpyDeviceToHost));
    });
}
extern "C" int32_t dcd_seq_op(dcd_ctx* c, int32_t op, const uint16_t* a, uint64_t na, int32_t Ra, const uint16_t* b, uint64_t nb,
                              int32_t Rb, uint16_t* out, uint64_t out_cap, uint64_t* out_n, int32_t* Rout) {
    if (!c || !out_n || (op != 0 && op != 1)) return DCD_E_INVALID;
    return guard2(c, [&] {
        cudaStream_t s = c->stream;
        const int absR = std::max(std::abs(Ra), std::abs(Rb));       // IndexSequence.cs:221-222
        const int R = (Ra < 0 && Rb < 0) ? -absR : absR;
        if (Rout) *Rout = R;
        const u64 cap = na + nb + 1;
        DevBuf<u16> da(c->pool, na + 1), db(c->pool, nb + 1);
        DevBuf<u64> va(c->pool, cap), vb(c->pool, cap), res(c->pool, cap), scr(c->pool, cap);
        u32 gallop_ratio = 16;
        if (const char* e = getenv("DCD_SEARCH_GALLOP")) gallop_ratio = (u32)std::max(0, atoi(e));
        DevBuf<u32> cnt(c->pool, 4);
        DevBuf<u8> tag(c->pool, cap);
        if (na) DCD_CUDA(cudaMemcpyAsync(da.p, a, na * 2, cudaMemcpyHostToDevice, s));
        if (nb) DCD_CUDA(cudaMemcpyAsync(db.p, b, nb * 2, cudaMemcpyHostToDevice, s));
        launch_seq_op(op, da.p, na, db.p, nb, (u32)absR, R < 0 ? 1u : 0u, va.p, vb.p, tag.p, scr.p, gallop_ratio, res.p, cnt.p, s);
        u32 m = 0;
        DCD_CUDA(cudaMemcpyAsync(&m, cnt.p, 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        // re-encode the result list (newSeq.AddRange, IndexSequence.cs:265,281) with the device encoder
        std::vector<u64> vals(m);
        if (m) DCD_CUDA(cudaMemcpy(vals.data(), res.p, (size_t)m * 8, cudaMemcpyDeviceToHost));
        int32_t st = dcd_seq_encode(c, reinterpret_cast<const uint64_t*>(vals.data()), m, out, out_cap, out_n);
        if (st != DCD_OK) throw StatusError(st, c->err);
    });
}

// ---- synthetic corpus -------------------------------------------------------------------------------
#include <cmath>

#include "synth.cuh"
extern "C" int32_t dcd_synth_generate(dcd_ctx* c, const uint8_t* words, const uint32_t* word_offs, int32_t nwords, double zipf_s,
                                      uint64_t seed, int64_t ndocs, int32_t pages_per_doc, int32_t min_page_chars, int64_t first_doc,
                                      dcd_synth* out) {
    if (!c || !out || nwords <= 0 || ndocs < 0 || pages_per_doc <= 0 || min_page_chars <= 0) return DCD_E_INVALID;
    return guard2(c, [&] {
        cudaStream_t s = c->stream;
        memset(out, 0, sizeof *out);
        std::vector<double> cdf((size_t)nwords);
        double h = 0;
        for (int i = 0; i < nwords; i++) h += std::pow((double)(i + 1), -zipf_s);
        double acc = 0;
        for (int i = 0; i < nwords; i++) { acc += std::pow((double)(i + 1), -zipf_s) / h; cdf[i] = acc; }
        cdf[nwords - 1] = 1.0;
        const u64 npages = (u64)ndocs * (u64)pages_per_doc;
        DevBuf<u8> d_words(c->pool, word_offs[nwords]);
        DevBuf<u32> d_woffs(c->pool, (size_t)nwords + 1);
        DevBuf<double> d_cdf(c->pool, (size_t)nwords);
        DevBuf<u64> page_len(c->pool, npages + 1), lb(c->pool, npages / 2048 + 2);
        DevBuf<u32> ticket(c->pool, 16);
        DCD_CUDA(cudaMemcpyAsync(d_words.p, words, word_offs[nwords], cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(d_woffs.p, word_offs, ((size_t)nwords + 1) * 4, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(d_cdf.p, cdf.data(), (size_t)nwords * 8, cudaMemcpyHostToDevice, s));
        SynthParams sp{d_words.p, d_woffs.p, d_cdf.p, (u32)nwords, seed, (u64)first_doc, (u64)ndocs, (u32)pages_per_doc, (u32)min_page_chars};
        u64* d_page_offs = (u64*)c->pool.alloc((npages + 1) * 8);
        u64* d_dfp = (u64*)c->pool.alloc(((u64)ndocs + 1) * 8);
        DCD_CUDA(cudaMemsetAsync(d_page_offs, 0, 8, s));
        DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
        launch_synth_lengths(sp, page_len.p, s);
        launch_inclusive_scan_u64(page_len.p, d_page_offs + 1, npages, 0, lb.p, ticket.p, s);
        u64 total = 0;
        if (npages) DCD_CUDA(cudaMemcpyAsync(&total, d_page_offs + npages, 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        u8* d_text = (u8*)c->pool.alloc(total + 64);
        launch_synth_write(sp, d_page_offs, d_text, s);
        launch_synth_doc_pages((u64)ndocs, (u32)pages_per_doc, d_dfp, s);
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        out->d_text = d_text;
        out->d_page_offs = reinterpret_cast<uint64_t*>(d_page_offs);
        out->d_doc_first_page = reinterpret_cast<uint64_t*>(d_dfp);
        out->text_bytes = total;
        out->npages = npages;
        out->ndocs = (u64)ndocs;
    });
}
extern "C" int32_t dcd_synth_free(dcd_ctx* c, dcd_synth* sy) {
    if (!c || !sy) return DCD_E_INVALID;
    if (sy->d_text) c->pool.release(sy->d_text);
    if (sy->d_page_offs) c->pool.release(sy->d_page_offs);
    if (sy->d_doc_first_page) c->pool.release(sy->d_doc_first_page);
    memset(sy, 0, sizeof *sy);
    return DCD_OK;
}
