// `.index` / `.index.list` byte formats (Build.cs:370-404, 99-148; Index.cs:312-380) and page-list metadata access.
#include <algorithm>
#include <cstring>
#include <numeric>

#include "engine.h"

namespace dcd {

__global__ void k_count_nonzero(const u64* __restrict__ v, u64 n, u64* __restrict__ out) {
    u64 c = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) c += v[i] != 0;
    c = warp_sum(c);
    if (lane_id() == 0 && c) atomicAdd(out, c);
}
u64 count_nonzero_u64(const u64* v, u64 n, u64* d_out, cudaStream_t s) {
    if (!n) return 0;
    unsigned g = (unsigned)std::min<u64>((n + 255) / 256, 148 * 8);
    k_count_nonzero<<<g, 256, 0, s>>>(v, n, d_out);
    u64 h = 0;
    DCD_CUDA(cudaMemcpyAsync(&h, d_out, 8, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    return h;
}

static void w_u64(std::vector<uint8_t>& o, u64 v) { for (int i = 0; i < 8; i++) o.push_back((uint8_t)(v >> (8 * i))); }
static void w_str(std::vector<uint8_t>& o, const u16s& s) {
    std::string u = utf16_to_utf8(s);
    u32 n = (u32)u.size();
    while (n >= 0x80) { o.push_back((uint8_t)(n | 0x80)); n >>= 7; }
    o.push_back((uint8_t)n);
    o.insert(o.end(), u.begin(), u.end());
}
static u16s dec(u64 v) {
    std::string t = std::to_string(v);
    return u16s(t.begin(), t.end());
}

void engine_fetch_page_ends(dcd_ctx_impl* c) {
    ResidentIndex& ix = c->ix;
    if (ix.h_page_end.size() == ix.npages) return;
    ix.h_page_end.resize(ix.npages);
    if (ix.npages) DCD_CUDA(cudaMemcpy(ix.h_page_end.data(), ix.d_page_end.p, ix.npages * 8, cudaMemcpyDeviceToHost));
}

void engine_fetch_segment_meta(PageSegment& sg) {
    if (!sg.bulk) return;
    if (sg.h_doc_first_page.size() != sg.ndocs + 1) {
        sg.h_doc_first_page.resize(sg.ndocs + 1);
        DCD_CUDA(cudaMemcpy(sg.h_doc_first_page.data(), sg.d_doc_first_page->p, (sg.ndocs + 1) * 8, cudaMemcpyDeviceToHost));
    }
    if (!sg.identity && sg.h_page_orig.size() != sg.count) {
        sg.h_page_orig.resize(sg.count);
        if (sg.count) DCD_CUDA(cudaMemcpy(sg.h_page_orig.data(), sg.d_page_orig->p, sg.count * 4, cudaMemcpyDeviceToHost));
    }
}

// (doc ordinal, id) of kept page k of a bulk segment
static void bulk_page_meta(PageSegment& sg, u64 k, u64* doc, u64* id_number) {
    engine_fetch_segment_meta(sg);
    const u64 orig = sg.identity ? k : sg.h_page_orig[k];
    // last d with h_doc_first_page[d] <= orig
    auto it = std::upper_bound(sg.h_doc_first_page.begin(), sg.h_doc_first_page.end(), orig);
    u64 d = (u64)(it - sg.h_doc_first_page.begin()) - 1;
    if (d >= sg.ndocs) d = sg.ndocs - 1;
    *doc = sg.first_doc + d;
    *id_number = orig - sg.h_doc_first_page[d] + 1;
}

u16s engine_doc_name(dcd_ctx_impl* c, u64 doc) {
    ResidentIndex& ix = c->ix;
    if (doc >= ix.ndocs) throw StatusError(DCD_E_INVALID, "document ordinal out of range");
    for (auto& sg : ix.segments)
        if (sg.bulk && doc >= sg.first_doc && doc < sg.first_doc + sg.ndocs) return sg.source + u":d" + dec(sg.doc_number_base + (doc - sg.first_doc));
    return doc < ix.doc_names.size() ? ix.doc_names[doc] : u16s();
}

void engine_page_info(dcd_ctx_impl* c, u64 ordinal, u16s* id, u64* doc, u64* start, u64* end) {
    ResidentIndex& ix = c->ix;
    if (!ix.valid) throw StatusError(DCD_E_STATE, "Index is not built");
    if (ordinal >= ix.npages) throw StatusError(DCD_E_INVALID, "page ordinal out of range");
    engine_fetch_page_ends(c);
    *end = ix.h_page_end[ordinal];
    *start = ordinal ? ix.h_page_end[ordinal - 1] : ix.coord_origin;   // IndexPageList.GetPage: page.coord = previous key (Build.cs:86-94)
    for (auto& sg : ix.segments) {
        if (ordinal < sg.first_ordinal || ordinal >= sg.first_ordinal + sg.count) continue;
        const u64 k = ordinal - sg.first_ordinal;
        if (sg.bulk) {
            u64 idn;
            bulk_page_meta(sg, k, doc, &idn);
            *id = dec(idn);
        } else {
            *doc = sg.page_doc[k];
            *id = sg.page_id[k];
        }
        return;
    }
    throw StatusError(DCD_E_INVALID, "page ordinal not covered by any segment");
}

// This rank's slice of the merged `.index` after a sharded build (Build.cs:376-386 layout; rank 0 carries the header).
void engine_serialize_merged(dcd_ctx_impl* c, std::vector<uint8_t>* io) {
    MergedIndex& mg = c->merged;
    if (!mg.valid) throw StatusError(DCD_E_STATE, "no merged index: call dcd_build_finish_sharded first");
    io->clear();
    const u32 n_own = mg.code_end - mg.code_begin;
    if (mg.h_off.size() != (size_t)n_own + 1) {
        mg.h_off.assign((size_t)n_own + 1, 0);
        if (mg.d_off.p) DCD_CUDA(cudaMemcpy(mg.h_off.data(), mg.d_off.p, ((size_t)n_own + 1) * 8, cudaMemcpyDeviceToHost));
    }
    std::vector<u16> arena(mg.units);
    if (mg.units) DCD_CUDA(cudaMemcpy(arena.data(), mg.d_arena.p, mg.units * 2, cudaMemcpyDeviceToHost));
    io->reserve(8 + mg.units * 2 + (size_t)n_own * 16);
    if (mg.rank == 0) w_u64(*io, mg.max_coord);
    for (u32 i = 0; i < n_own; i++) {
        const u64 a = mg.h_off[i], e = mg.h_off[i + 1];
        const int32_t n = (int32_t)(e - a);
        w_str(*io, c->code_string(mg.code_begin + i));
        for (int k = 0; k < 4; k++) io->push_back((uint8_t)((u32)n >> (8 * k)));
        const uint8_t* p = (const uint8_t*)(arena.data() + a);
        io->insert(io->end(), p, p + (size_t)n * 2);
    }
}

// Builder.Save (Build.cs:370-404) + IndexPageList.Save (Build.cs:99-118)
void engine_serialize(dcd_ctx_impl* c, std::vector<uint8_t>* io, std::vector<uint8_t>* lo) {
    ResidentIndex& ix = c->ix;
    if (!ix.valid) throw StatusError(DCD_E_STATE, "Index is not built");
    io->clear();
    lo->clear();
    std::vector<u16> arena(ix.arena_units);
    if (ix.arena_units) DCD_CUDA(cudaMemcpy(arena.data(), ix.d_arena.p, ix.arena_units * 2, cudaMemcpyDeviceToHost));
    io->reserve(8 + ix.arena_units * 2 + (size_t)ix.ncodes * 16);
    w_u64(*io, ix.max_coord);
    for (u32 k = 0; k < ix.ncodes; k++) {
        const u64 a = ix.h_code_off[k], e = ix.h_code_off[k + 1];
        const int32_t n = (int32_t)(e - a);
        if (!n && ix.global_codes) continue;   // key of another shard (sharded build): this shard has no list for it
        w_str(*io, c->code_string(k));
        for (int i = 0; i < 4; i++) io->push_back((uint8_t)((u32)n >> (8 * i)));
        const uint8_t* p = (const uint8_t*)(arena.data() + a);
        io->insert(io->end(), p, p + (size_t)n * 2);
    }
    engine_fetch_page_ends(c);
    i64 last_doc = -1;
    for (auto& sg : ix.segments) {
        if (sg.bulk) engine_fetch_segment_meta(sg);
        u64 d = 0;   // running doc pointer inside a bulk segment
        for (u64 k = 0; k < sg.count; k++) {
            const u64 end = ix.h_page_end[sg.first_ordinal + k];
            u64 doc;
            u16s id;
            if (sg.bulk) {
                const u64 orig = sg.identity ? k : sg.h_page_orig[k];
                while (d + 1 < sg.ndocs && sg.h_doc_first_page[d + 1] <= orig) d++;
                doc = sg.first_doc + d;
                id = dec(orig - sg.h_doc_first_page[d] + 1);
            } else {
                doc = sg.page_doc[k];
                id = sg.page_id[k];
            }
            w_u64(*lo, end);
            if ((i64)doc != last_doc) {
                last_doc = (i64)doc;
                w_str(*lo, engine_doc_name(c, doc));
                w_u64(*lo, end);
            }
            w_str(*lo, u":" + id);
        }
    }
}

struct Rd {
    const uint8_t* p;
    size_t n, i = 0;
    bool eof = false;
    u64 r_u64() {
        if (i + 8 > n) { eof = true; return 0; }
        u64 v = 0;
        for (int k = 0; k < 8; k++) v |= (u64)p[i + k] << (8 * k);
        i += 8;
        return v;
    }
    int32_t r_i32() {
        if (i + 4 > n) { eof = true; return 0; }
        u32 v = 0;
        for (int k = 0; k < 4; k++) v |= (u32)p[i + k] << (8 * k);
        i += 4;
        return (int32_t)v;
    }
    u16s r_str() {
        u32 len = 0;
        int sh = 0;
        while (true) {
            if (i >= n) { eof = true; return u16s(); }
            uint8_t b = p[i++];
            len |= (u32)(b & 0x7F) << sh;
            sh += 7;
            if (!(b & 0x80)) break;
        }
        if (i + len > n) { eof = true; return u16s(); }
        u16s s = utf8_to_utf16(p + i, len);
        i += len;
        return s;
    }
};

// Index.Load (Index.cs:312-380) + IndexPageList.Load (Build.cs:119-148): makes an index written by stock Docodo
// (or by engine_serialize) resident on the device.
void engine_load_bytes(dcd_ctx_impl* c, const uint8_t* ib, u64 ni, const uint8_t* lb, u64 nl) {
    ResidentIndex& ix = c->ix;
    ix.clear();
    Rd r{ib, (size_t)ni};
    ix.max_coord = r.r_u64();
    if (r.eof) throw StatusError(DCD_E_FORMAT, ".index is shorter than its header");
    std::vector<u16s> keys;
    std::vector<std::pair<size_t, size_t>> spans;   // byte offset, ushort count
    while (true) {
        u16s k = r.r_str();
        if (r.eof) break;
        int32_t n = r.r_i32();
        if (r.eof || n < 0 || r.i + (size_t)n * 2 > r.n) break;
        keys.push_back(k);
        spans.emplace_back(r.i, (size_t)n);
        r.i += (size_t)n * 2;
    }
    // SortedList with the ordinal comparer (Index.cs:142-151)
    std::vector<u32> order(keys.size());
    std::iota(order.begin(), order.end(), 0u);
    if (!std::is_sorted(keys.begin(), keys.end()))
        std::stable_sort(order.begin(), order.end(), [&](u32 a, u32 b) { return keys[a] < keys[b]; });
    ix.ncodes = (u32)keys.size();
    ix.h_code_off.assign(keys.size() + 1, 0);
    ix.h_code_str_off.assign(keys.size() + 1, 0);
    std::vector<u16> arena;
    for (size_t q = 0; q < order.size(); q++) {
        const u32 k = order[q];
        ix.h_code_chars.append((const u16*)keys[k].data(), keys[k].size());
        ix.h_code_str_off[q + 1] = (u32)ix.h_code_chars.size();
        const size_t at = arena.size();
        arena.resize(at + spans[k].second);
        if (spans[k].second) memcpy(arena.data() + at, ib + spans[k].first, spans[k].second * 2);
        ix.h_code_off[q + 1] = arena.size();
    }
    ix.arena_units = arena.size();
    ix.d_arena = DevBuf<u16>(c->pool, arena.size());
    ix.d_code_off = DevBuf<u64>(c->pool, keys.size() + 1);
    if (!arena.empty()) DCD_CUDA(cudaMemcpy(ix.d_arena.p, arena.data(), arena.size() * 2, cudaMemcpyHostToDevice));
    DCD_CUDA(cudaMemcpy(ix.d_code_off.p, ix.h_code_off.data(), (keys.size() + 1) * 8, cudaMemcpyHostToDevice));
    // page list
    Rd l{lb, (size_t)nl};
    PageSegment sg;
    sg.bulk = false;
    while (true) {
        u64 n = l.r_u64();
        if (l.eof) break;
        u16s s = l.r_str();
        if (l.eof || s.empty()) break;
        if (s[0] != u':') {
            ix.doc_names.push_back(s);
        } else {
            if (ix.doc_names.empty()) ix.doc_names.push_back(u16s());
            ix.h_page_end.push_back(n);
            sg.page_doc.push_back((u32)ix.doc_names.size() - 1);
            sg.page_id.push_back(s.substr(1));
        }
    }
    sg.count = ix.h_page_end.size();
    sg.first_ordinal = 0;
    ix.npages = sg.count;
    ix.ndocs = ix.doc_names.size();
    ix.segments.push_back(std::move(sg));
    ix.d_page_end = DevBuf<u64>(c->pool, ix.npages);
    if (ix.npages) DCD_CUDA(cudaMemcpy(ix.d_page_end.p, ix.h_page_end.data(), ix.npages * 8, cudaMemcpyHostToDevice));
    ix.valid = true;
    c->counters.codes = ix.ncodes;
    c->counters.encoded_u16 = ix.arena_units;
    c->counters.max_coord = ix.max_coord;
    c->counters.pages = ix.npages;
}

}  // namespace dcd
