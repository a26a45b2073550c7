// Host-side engine behind the C ABI (include/docodo_b200.h): context, device memory pool, build session,
// resident index.  Mirrors the state of Docodo.Index / Index.Builder (Index.cs, Build.cs).
#pragma once
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/docodo_b200.h"
#include "build_kernels.cuh"
#include "common.cuh"
#include "strtable.cuh"

namespace dcd {

using u16s = std::u16string;

// Caching device allocator: benchmark steps re-use the previous step's blocks instead of paying cudaMalloc/cudaFree.
// Memory is managed as segments of cudaMalloc'ed slabs: a live segment can be SHRUNK in place (the tail becomes a free
// segment) -- the encoded index adopts the front of a sort buffer without a copy -- and free neighbours coalesce.
struct DevPool {
    struct Seg { size_t size; bool free; char* slab; };
    std::map<char*, Seg> segs;                       // by address
    std::multimap<size_t, char*> free_by_size;
    std::unordered_map<char*, size_t> slabs;         // slab base -> size
    size_t bytes_live = 0, bytes_cached = 0, peak = 0;
    void* alloc(size_t bytes);
    void release(void* p);
    void shrink(void* p, size_t new_bytes);          // keep the first new_bytes of a live block
    void trim();                                     // cudaFree every slab that is entirely free
    ~DevPool();
private:
    void add_free(char* p);
    void remove_free(char* p);
};

template <typename T>
struct DevBuf {   // RAII view on a pool block
    DevPool* pool = nullptr;
    T* p = nullptr;
    size_t n = 0;
    DevBuf() {}
    DevBuf(DevPool& pl, size_t count) : pool(&pl), n(count) { p = (T*)pl.alloc(std::max<size_t>(count, 1) * sizeof(T)); }
    DevBuf(DevBuf&& o) noexcept : pool(o.pool), p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        reset();
        pool = o.pool; p = o.p; n = o.n;
        o.p = nullptr; o.n = 0;
        return *this;
    }
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    void reset() { if (p && pool) pool->release(p); p = nullptr; n = 0; }
    // take over the block of `o` (any element type), keeping only its first `count` elements
    template <typename U>
    void adopt(DevBuf<U>&& o, size_t count) {
        reset();
        pool = o.pool; p = reinterpret_cast<T*>(o.p); n = count;
        o.p = nullptr; o.n = 0;
        if (p && pool) pool->shrink(p, std::max<size_t>(count, 1) * sizeof(T));
    }
    ~DevBuf() { reset(); }
    size_t bytes() const { return n * sizeof(T); }
};

// Page-locked host array for the device -> host copies of the build (key table) so that they run at PCIe speed and
// asynchronously.  Blocks are recycled through a process-wide cache: benchmark steps do not pay cudaMallocHost again.
void* pinned_cache_alloc(size_t bytes, size_t* cap_out);
void pinned_cache_release(void* p, size_t cap);
template <typename T>
struct PinnedVec {
    T* p = nullptr;
    size_t n = 0, cap_bytes = 0;
    PinnedVec() {}
    PinnedVec(const PinnedVec&) = delete;
    PinnedVec& operator=(const PinnedVec&) = delete;
    PinnedVec(PinnedVec&& o) noexcept : p(o.p), n(o.n), cap_bytes(o.cap_bytes) { o.p = nullptr; o.n = 0; o.cap_bytes = 0; }
    PinnedVec& operator=(PinnedVec&& o) noexcept {
        if (this != &o) { drop(); p = o.p; n = o.n; cap_bytes = o.cap_bytes; o.p = nullptr; o.n = 0; o.cap_bytes = 0; }
        return *this;
    }
    ~PinnedVec() { drop(); }
    void drop() { if (p) pinned_cache_release(p, cap_bytes); p = nullptr; n = 0; cap_bytes = 0; }
    void reserve(size_t count) {
        if (count * sizeof(T) <= cap_bytes) return;
        size_t nc = 0;
        T* q = (T*)pinned_cache_alloc(std::max(count * sizeof(T), cap_bytes + cap_bytes / 2), &nc);
        if (n) memcpy(q, p, n * sizeof(T));
        if (p) pinned_cache_release(p, cap_bytes);
        p = q; cap_bytes = nc;
    }
    void resize_uninit(size_t count) { reserve(count); n = count; }   // contents undefined: the caller overwrites them
    void assign(size_t count, T v) { resize_uninit(count); for (size_t i = 0; i < count; i++) p[i] = v; }
    void append(const T* src, size_t count) { reserve(n + count); memcpy(p + n, src, count * sizeof(T)); n += count; }
    T* data() { return p; }
    const T* data() const { return p; }
    size_t size() const { return n; }
    T& operator[](size_t i) { return p[i]; }
    const T& operator[](size_t i) const { return p[i]; }
};

struct StrTableOwner {   // device arrays of one StrTable
    DevBuf<u64> slot_hash;
    DevBuf<u32> slot_entry, ent_off, ent_len, ent_flags, ent_value, counters;
    DevBuf<u16> arena;
    StrTable view{};
    void create(DevPool& pool, u32 cap_log2, u32 max_entries, u32 arena_units, cudaStream_t s);
    void clear(cudaStream_t s);
};

struct VocabHost {   // Vocab (Dict.cs:13-104)
    std::unordered_map<u16s, int32_t> map;
    std::vector<std::pair<u16s, int32_t>> ordered;   // enumeration order of the reference's Dictionary == insertion (file) order
    char16_t begin = 0, end = 0;
    int stemmer = 0;
    StrTableOwner dev;
};

// The page list is a sequence of segments in indexing order.  Explicit segments carry per-page records (ids, doc
// ordinals); bulk segments have implicit metadata (ids "1".."k" inside each document, docs "<source>:d<number>").
struct PageSegment {
    bool bulk = false;
    u64 first_ordinal = 0;     // position of the segment's first kept page in the global page list
    u64 count = 0;             // kept pages in the segment
    // explicit
    std::vector<u32> page_doc; // global doc ordinal per page
    std::vector<u16s> page_id;
    // bulk
    u64 first_doc = 0;         // global ordinal of the segment's first document
    u64 ndocs = 0;
    u64 doc_number_base = 0;   // "d<doc_number_base + k>"
    u16s source;
    u64 npages_orig = 0;
    bool identity = true;      // no empty page was dropped: kept page k == original page k
    std::vector<u64> h_doc_first_page;   // [ndocs + 1], original page numbering (lazily downloaded)
    std::vector<u32> h_page_orig;        // kept -> original page index when !identity (lazily downloaded)
    std::shared_ptr<DevBuf<u64>> d_doc_first_page;
    std::shared_ptr<DevBuf<u32>> d_page_orig;
};

struct ResidentIndex {   // what Index.Load leaves in memory (Index.cs:312-380), on the device
    bool valid = false;
    bool global_codes = false;        // sharded build: the key table is the union over ranks, local lists may be empty
    DevBuf<u64> d_code_base;          // sharded build (> 1 rank): per code, the coordinate the list's first delta is taken against
                                      // (last coordinate of the code on the earlier shards); empty otherwise
    DevBuf<u64> d_next_first;         // exact sharded serving: per code, its first coordinate on the LATER shards (~0: none); filled
    int next_first_ranks = 0;         // by the first sharded batch (search.cu, ensure_next_first) together with every shard's origin
    std::vector<u64> shard_origins;
    u64 max_coord = 0;
    u64 coord_origin = 0;             // coordinate of the first character (document-range shards, dcd_build_set_origin)
    u32 ncodes = 0;
    u64 arena_units = 0;
    DevBuf<u16> d_arena;              // all IndexSequences back to back, code order
    DevBuf<u64> d_code_off;           // [ncodes + 1]
    PinnedVec<u64> h_code_off;        // host copy
    // decoded-list cache (search.cu): the long lists the query batches keep asking for, decoded once per resident index
    std::unordered_map<u32, std::pair<u64*, u32>> dcache;   // code -> (device list, values)
    std::vector<DevBuf<u64>> dcache_bufs;
    u64 dcache_bytes = 0;
    std::vector<u32> h_code_nvals;    // coordinates per code (terminator words), made by the first batch that needs it (search.cu)
    PinnedVec<u16> h_code_chars;      // code strings, ordinal order
    PinnedVec<u32> h_code_str_off;    // [ncodes + 1]
    DevBuf<u16> d_key_chars;          // device copy of the key strings for the dictionary-side scans (dict_kernels.cu), made on
    DevBuf<u32> d_key_off;            // first use
    // page list
    u64 npages = 0;
    DevBuf<u64> d_page_end;           // [npages] ascending end coordinates of the kept pages
    std::vector<u64> h_page_end;      // lazily filled
    std::vector<PageSegment> segments;
    std::vector<u16s> doc_names;      // explicit documents by global ordinal (empty string for bulk docs)
    u64 ndocs = 0;
    std::vector<u16> seq_scratch;     // dcd_index_sequence result
    u16s tmp_id, tmp_name;
    void clear() { *this = ResidentIndex(); }
    ResidentIndex() = default;
    ResidentIndex(ResidentIndex&&) = default;
    ResidentIndex& operator=(ResidentIndex&&) = default;
};

// This rank's code range of the MERGED index after a sharded build (shard.h): the concatenation over ranks, in rank
// order, of arena[code_off[0] .. code_off[code_end - code_begin]) is the single-builder `.index` payload.
struct MergedIndex {
    bool valid = false;
    int rank = 0, nranks = 1;
    u32 code_begin = 0, code_end = 0;
    u64 units = 0;
    u64 max_coord = 0;                // global Index.MaxCoord (largest over ranks)
    DevBuf<u16> d_arena;
    DevBuf<u64> d_off;                // [code_end - code_begin + 1]
    std::vector<u64> h_off;           // lazily downloaded
    void clear() { *this = MergedIndex(); }
    MergedIndex() = default;
    MergedIndex(MergedIndex&&) = default;
    MergedIndex& operator=(MergedIndex&&) = default;
};

struct ShardComm;

struct dcd_ctx_impl {
    int device = 0;
    int nsm = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // device -> host copies that overlap the build kernels
    cudaEvent_t copy_event = nullptr;
    cudaStream_t ce_stream[4] = {};       // sharded build: copy-engine transfers of the big segments
    cudaStream_t xchg_stream = nullptr;   // sharded build: the scatter of one owner's slice runs here while the next slice is encoded
    cudaEvent_t xchg_event[65] = {};      // piece encoded (0..63), scatter finished (64)
    bool own_stream = true;    // false after dcd_ctx_set_stream: the caller's stream is borrowed
    std::string err;
    dcd_options opt{};
    DevPool pool;
    std::shared_ptr<void> pages_sink;   // dcd_search_batch_pages .. dcd_search_pages_fetch: PagesSink (search.h), device-resident pairs
    // unicode tables
    std::vector<u8> h_cls;
    std::vector<u16> h_lower, h_upper;
    DevBuf<u8> d_cls;
    DevBuf<u16> d_lower;
    // Index configuration
    std::vector<std::unique_ptr<VocabHost>> vocs;
    std::unordered_set<u16s> stop;
    ResidentIndex ix;
    MergedIndex merged;
    dcd_counters counters{};
    // helpers
    u16s lower(const u16s& s) const;
    u16s upper(const u16s& s) const;
    bool is_w(char16_t c) const { return h_cls[(u16)c] & CLS_W; }
    bool is_L(char16_t c) const { return h_cls[(u16)c] & CLS_L; }
    std::vector<u16s> word_codes(const u16s& word) const;   // GetWordCodes on the host (query terms)
    int find_code(const u16s& code) const;                  // ordinal code id or -1
    u16s code_string(u32 id) const;
};

struct TextBatch {
    DevBuf<u8> own_text;       // owned copy (host input) or empty (caller-owned device pointer)
    const void* text = nullptr;
    int enc = 0;
    u64 n_units = 0;
    u64 npages = 0;
    DevBuf<u64> own_page_off;
    const u64* page_off = nullptr;
    DevBuf<i64> page_delta;
    DevBuf<u64> coord_end;     // bulk: per original page
    DevBuf<u32> tile_first_page;
    u32 ntiles = 0;
    int segment = -1;          // bulk: index into build->segments
    bool uniform = false;      // every page has the same coordinate delta (bulk pages): the v2 tokenizer applies
    i64 delta0 = 0;            // that delta: coord = unit offset + delta0
    u32 ntiles2 = 0;           // v2 tiles (960 units)
    // host input: the text travels in chunks on the context's copy stream; (units resident, event) per chunk, so that
    // the tokenizer starts on the first chunks while the rest is still on the PCIe bus.  Declared last: destroyed
    // (= copies drained) before own_text goes back to the pool.
    struct H2DChunks {
        std::vector<std::pair<u64, cudaEvent_t>> ev;
        ~H2DChunks() {
            for (auto& e : ev) { cudaEventSynchronize(e.second); cudaEventDestroy(e.second); }
        }
    };
    std::shared_ptr<H2DChunks> h2d;
};

struct dcd_build_impl {
    dcd_ctx_impl* ctx = nullptr;
    u64 coord = 0;             // IndexTask local `coord` (Build.cs:459)
    u64 coord_origin = 0;      // dcd_build_set_origin
    double add_ms = 0;         // host wall time spent in dcd_build_add_pages_bulk
    u64 max_coord = 0;         // Builder.maxCoord (Build.cs:294)
    bool any_posting = false;
    // page list under construction (segments in indexing order)
    std::vector<PageSegment> segments;
    std::vector<u16s> doc_names;        // global doc ordinal -> name for explicit docs
    std::vector<u64> explicit_page_end; // end coordinate per explicit page, in order (all explicit segments)
    u64 ndocs = 0;
    u64 bulk_docs_named = 0;
    // host-staged body text of explicit pages
    std::vector<u16> host_text;
    std::vector<u64> host_page_off;     // offsets into host_text
    std::vector<i64> host_page_delta;
    // host-supplied postings (header pages, Builder API)
    std::vector<u16> extra_chars;
    std::vector<u32> extra_offs{0};
    std::vector<u64> extra_coord;
    std::vector<u8> extra_kind;         // 0 AddWord, 1 Add
    std::vector<TextBatch> batches;
};

// comm == nullptr: single-builder build.  Otherwise the collective sharded build of shard.h: every rank of `comm` calls
// it at the same time; the resident index keeps this rank's document range (global coordinates, global code ids) and
// ctx->merged receives this rank's code range of the merged index.
void engine_build_finish(dcd_build_impl* b, ShardComm* comm = nullptr);
void engine_serialize(dcd_ctx_impl* c, std::vector<uint8_t>* index_out, std::vector<uint8_t>* list_out);
void engine_serialize_merged(dcd_ctx_impl* c, std::vector<uint8_t>* index_out);
void engine_load_bytes(dcd_ctx_impl* c, const uint8_t* ib, u64 ni, const uint8_t* lb, u64 nl);
void engine_page_info(dcd_ctx_impl* c, u64 ordinal, u16s* id, u64* doc, u64* start, u64* end);
u16s engine_doc_name(dcd_ctx_impl* c, u64 doc);
void engine_fetch_page_ends(dcd_ctx_impl* c);
void engine_fetch_segment_meta(PageSegment& sg);   // host copies of a bulk segment's document / page maps

// Dictionary-side queries over the resident key table, dict_kernels.cu
std::vector<u16s> engine_like_words(dcd_ctx_impl* c, const u16s& word);                  // GetLikeWords, Search.cs:160-167
std::vector<u16s> engine_close_words(dcd_ctx_impl* c, const u16s& word);                 // GetCloseWords, Search.cs:169-174
std::vector<u16s> engine_suggestions(dcd_ctx_impl* c, const u16s& req, int n);           // GetSuggessions, Search.cs:176-188
std::vector<std::pair<u16s, u64>> engine_histogram(dcd_ctx_impl* c, int top);            // CalcHistogram, Index.cs:284-307

std::string utf16_to_utf8(const u16s& s);
u16s utf8_to_utf16(const uint8_t* p, size_t n);

}  // namespace dcd

struct dcd_ctx : dcd::dcd_ctx_impl {};
struct dcd_comm { dcd::ShardComm* impl = nullptr; };
struct dcd_build : dcd::dcd_build_impl {};
