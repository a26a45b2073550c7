// extern "C" entry points of libdocodo_b200.so (include/docodo_b200.h).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <mutex>

#include "encode.cuh"
#include "engine.h"
#include "utf8.cuh"
#include "shard.h"
#include "search.h"
#include "sort.cuh"
#include "stem_en.cuh"
#include "stem_ru.cuh"
#include "synth.cuh"
#include "unicode_tables.h"

using namespace dcd;


static std::string g_last_error;
static std::mutex g_err_mu;

template <typename F>
static int32_t guard(dcd_ctx_impl* c, F&& f) {
    try {
        if (c) DCD_CUDA(cudaSetDevice(c->device));
        f();
        return DCD_OK;
    } catch (const StatusError& e) {
        if (c) c->err = e.what(); else { std::lock_guard<std::mutex> l(g_err_mu); g_last_error = e.what(); }
        return e.code;
    } catch (const CudaError& e) {
        cudaGetLastError();
        if (c) c->err = e.what(); else { std::lock_guard<std::mutex> l(g_err_mu); g_last_error = e.what(); }
        return DCD_E_CUDA;
    } catch (const std::exception& e) {
        if (c) c->err = e.what(); else { std::lock_guard<std::mutex> l(g_err_mu); g_last_error = e.what(); }
        return DCD_E_INVALID;
    }
}

namespace dcd {

u16s dcd_ctx_impl::lower(const u16s& s) const {
    u16s r = s;
    for (auto& ch : r) ch = (char16_t)h_lower[(u16)ch];
    return r;
}
u16s dcd_ctx_impl::upper(const u16s& s) const {
    u16s r = s;
    for (auto& ch : r) ch = (char16_t)h_upper[(u16)ch];
    return r;
}

// Host evaluation of GetWordCodes (Build.cs:175-247) for query terms and the Builder API, using the same
// host+device stemmer sources as the K4 kernel.
static u16s host_stem(int kind, const u16s& w) {
    if (kind != DCD_STEM_EN && kind != DCD_STEM_RU) return w;
    std::vector<u16> buf(w.size() + 4);
    for (size_t i = 0; i < w.size(); i++) buf[i] = (u16)w[i];
    int n = kind == DCD_STEM_EN ? en::stem(buf.data(), (int)w.size()) : ru::stem(buf.data(), (int)w.size());
    return u16s((const char16_t*)buf.data(), n);
}
static bool host_in_range(const u16s& w, int k) {
    for (char16_t ch : w) {
        u16 c = (u16)ch;
        bool ok;
        switch (k) {
            case 0: ok = c >= '0' && c <= '9'; break;
            case 1: ok = (c >= 0x0430 && c <= 0x044F) || c == 0x0451; break;
            case 2: ok = c >= 'a' && c <= 'z'; break;
            case 3: ok = (c >= 'a' && c <= 'z') || c == 0x1E9E || c == 0xE4 || c == 0xFC || c == 0xF6; break;
            default:
                ok = (c >= 'a' && c <= 'z') || c == 0xE9 || c == 0xE2 || c == 0xE0 || c == 0xEA || c == 0xE8 || c == 0xEB ||
                     c == 0xE7 || c == 0xEE || c == 0xEF || c == 0xF4 || c == 0xFB || c == 0xF9 || c == 0xFC || c == 0xFF;
        }
        if (!ok) return false;
    }
    return true;
}
std::vector<u16s> dcd_ctx_impl::word_codes(const u16s& word) const {
    std::vector<u16s> l;
    if (word.empty()) return l;
    if (word[0] >= u'0' && word[0] <= u'9') { l.push_back(word); return l; }
    if (stop.count(word)) return l;
    l.push_back(word);
    u16s stemmed = word, first;
    int nG = 0;
    for (size_t v = 0; v < vocs.size(); v++) {
        const VocabHost& voc = *vocs[v];
        if (word[0] >= voc.begin && word[0] <= voc.end) {
            stemmed = host_stem(voc.stemmer, word);
            auto it = voc.map.find(stemmed);
            nG = it == voc.map.end() ? 0 : it->second;
            if (nG != 0) {
                char buf[16];
                snprintf(buf, sizeof buf, "%X", (unsigned)(((u32)v << 24) | ((u32)nG & 0xFFFFFFu)));
                u16s code = u"#";
                for (char* p = buf; *p; p++) code.push_back((char16_t)*p);
                l.push_back(code);
            }
        }
        if (first.empty()) first = stemmed;
    }
    if (nG == 0) {
        stemmed = first;
        if (vocs.empty()) {
            for (int k = 0; k < 5; k++) {
                if (!(opt.stemmers_mask & (1 << k))) continue;
                if (host_in_range(word, k)) {
                    if (k == 1) stemmed = host_stem(DCD_STEM_RU, word);
                    else if (k == 2) stemmed = host_stem(DCD_STEM_EN, word);
                    break;
                }
            }
        }
        if (!stemmed.empty() && stemmed != word) l.push_back(u"$" + stemmed);
    }
    return l;
}

int dcd_ctx_impl::find_code(const u16s& code) const {
    // binary search over the ordinal-sorted packed code strings (SortedList.Keys.Contains, Search.cs:235)
    i64 lo = 0, hi = (i64)ix.ncodes - 1;
    const u16* chars = ix.h_code_chars.data();
    while (lo <= hi) {
        i64 mid = (lo + hi) >> 1;
        const u16* s = chars + ix.h_code_str_off[mid];
        size_t n = ix.h_code_str_off[mid + 1] - ix.h_code_str_off[mid];
        size_t m = std::min(n, code.size());
        int cmp = 0;
        for (size_t i = 0; i < m; i++) {
            if (s[i] != (u16)code[i]) { cmp = s[i] < (u16)code[i] ? -1 : 1; break; }
        }
        if (cmp == 0) cmp = n < code.size() ? -1 : (n > code.size() ? 1 : 0);
        if (cmp == 0) return (int)mid;
        if (cmp < 0) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}
u16s dcd_ctx_impl::code_string(u32 id) const {
    return u16s((const char16_t*)ix.h_code_chars.data() + ix.h_code_str_off[id], ix.h_code_str_off[id + 1] - ix.h_code_str_off[id]);
}

std::string utf16_to_utf8(const u16s& s) {   // BinaryWriter.Write(string) payload; lone surrogates -> U+FFFD like UTF8Encoding
    std::string o;
    for (size_t i = 0; i < s.size(); i++) {
        u32 c = (u16)s[i];
        if (c >= 0xD800 && c <= 0xDBFF && i + 1 < s.size() && (u16)s[i + 1] >= 0xDC00 && (u16)s[i + 1] <= 0xDFFF) {
            c = 0x10000 + ((c - 0xD800) << 10) + ((u16)s[i + 1] - 0xDC00);
            i++;
        } else if (c >= 0xD800 && c <= 0xDFFF) {
            c = 0xFFFD;
        }
        if (c < 0x80) o.push_back((char)c);
        else if (c < 0x800) { o.push_back((char)(0xC0 | (c >> 6))); o.push_back((char)(0x80 | (c & 0x3F))); }
        else if (c < 0x10000) {
            o.push_back((char)(0xE0 | (c >> 12)));
            o.push_back((char)(0x80 | ((c >> 6) & 0x3F)));
            o.push_back((char)(0x80 | (c & 0x3F)));
        } else {
            o.push_back((char)(0xF0 | (c >> 18)));
            o.push_back((char)(0x80 | ((c >> 12) & 0x3F)));
            o.push_back((char)(0x80 | ((c >> 6) & 0x3F)));
            o.push_back((char)(0x80 | (c & 0x3F)));
        }
    }
    return o;
}
u16s utf8_to_utf16(const uint8_t* p, size_t n) {
    u16s o;
    size_t i = 0;
    while (i < n) {
        u32 c = p[i];
        int extra;
        if (c < 0x80) extra = 0;
        else if ((c & 0xE0) == 0xC0) { c &= 0x1F; extra = 1; }
        else if ((c & 0xF0) == 0xE0) { c &= 0x0F; extra = 2; }
        else if ((c & 0xF8) == 0xF0) { c &= 0x07; extra = 3; }
        else { o.push_back(0xFFFD); i++; continue; }
        if (i + extra >= n && extra) { o.push_back(0xFFFD); i++; continue; }
        bool ok = true;
        for (int k = 1; k <= extra; k++) {
            if ((p[i + k] & 0xC0) != 0x80) { ok = false; break; }
            c = (c << 6) | (p[i + k] & 0x3F);
        }
        if (!ok) { o.push_back(0xFFFD); i++; continue; }
        i += extra + 1;
        if (c >= 0x10000) { c -= 0x10000; o.push_back((char16_t)(0xD800 + (c >> 10))); o.push_back((char16_t)(0xDC00 + (c & 0x3FF))); }
        else o.push_back((char16_t)c);
    }
    return o;
}

}  // namespace dcd

namespace dcd { u64 count_nonzero_u64(const u64*, u64, u64*, cudaStream_t); }
static u16s mk(const uint16_t* p, size_t n) { return u16s((const char16_t*)p, n); }

// ===================================================================================================
// context
// ===================================================================================================
extern "C" int32_t dcd_options_default(dcd_options* o) {
    if (!o) return DCD_E_INVALID;
    memset(o, 0, sizeof *o);
    o->min_word_length = 3;
    o->max_word_length = 32;
    o->max_found_pages = 30000;
    o->max_found_docs = 500;
    o->keep_forms = 1;
    o->stemmers_mask = 0x1F;
    o->dict_capacity_log2 = 23;
    return DCD_OK;
}

extern "C" int32_t dcd_ctx_create(int32_t device, dcd_ctx** out) {
    if (!out) return DCD_E_INVALID;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0 || device < 0 || device >= count) {
        cudaGetLastError();
        std::lock_guard<std::mutex> l(g_err_mu);
        g_last_error = "no usable CUDA device (docodo_b200 has no CPU fallback)";
        return DCD_E_NO_DEVICE;
    }
    cudaDeviceProp prop{};
    cudaGetDeviceProperties(&prop, device);
    if (prop.major < 10) {
        std::lock_guard<std::mutex> l(g_err_mu);
        g_last_error = std::string("device ") + prop.name + " is not sm_100: libdocodo_b200 is built for sm_100a only";
        return DCD_E_NO_DEVICE;
    }
    dcd_ctx* c = new dcd_ctx();
    int32_t st = guard(nullptr, [&] {
        c->device = device;
        DCD_CUDA(cudaSetDevice(device));
        c->nsm = prop.multiProcessorCount;
        DCD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        DCD_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        DCD_CUDA(cudaEventCreateWithFlags(&c->copy_event, cudaEventDisableTiming));
        DCD_CUDA(cudaStreamCreateWithFlags(&c->xchg_stream, cudaStreamNonBlocking));
        for (auto& st : c->ce_stream) DCD_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        for (auto& e : c->xchg_event) DCD_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        dcd_options_default(&c->opt);
        c->h_cls.assign(65536, 0);
        c->h_lower.resize(65536);
        c->h_upper.resize(65536);
        for (int i = 0; i < DCD_N_CLS_RUNS; i++)
            for (u32 ch = DCD_CLS_RUNS[i][0]; ch <= DCD_CLS_RUNS[i][1]; ch++) c->h_cls[ch] = (u8)DCD_CLS_RUNS[i][2];
        for (u32 ch = 0; ch < 65536; ch++) c->h_lower[ch] = c->h_upper[ch] = (u16)ch;
        for (int i = 0; i < DCD_N_LOWER_PAIRS; i++) c->h_lower[DCD_LOWER_PAIRS[i][0]] = DCD_LOWER_PAIRS[i][1];
        for (int i = 0; i < DCD_N_UPPER_PAIRS; i++) c->h_upper[DCD_UPPER_PAIRS[i][0]] = DCD_UPPER_PAIRS[i][1];
        c->d_cls = DevBuf<u8>(c->pool, 65536);
        c->d_lower = DevBuf<u16>(c->pool, 65536);
        DCD_CUDA(cudaMemcpy(c->d_cls.p, c->h_cls.data(), 65536, cudaMemcpyHostToDevice));
        DCD_CUDA(cudaMemcpy(c->d_lower.p, c->h_lower.data(), 65536 * 2, cudaMemcpyHostToDevice));
    });
    if (st != DCD_OK) { delete c; return st; }
    *out = c;
    return DCD_OK;
}

extern "C" void dcd_ctx_destroy(dcd_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->pages_sink.reset();
    c->ix.clear();
    c->vocs.clear();
    c->d_cls.reset();
    c->d_lower.reset();
    if (c->own_stream) cudaStreamDestroy(c->stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->copy_event) cudaEventDestroy(c->copy_event);
    if (c->xchg_stream) cudaStreamDestroy(c->xchg_stream);
    for (auto st : c->ce_stream) if (st) cudaStreamDestroy(st);
    for (auto e : c->xchg_event) if (e) cudaEventDestroy(e);
    delete c;
}

extern "C" int32_t dcd_ctx_set_stream(dcd_ctx* c, void* cuda_stream) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        DCD_CUDA(cudaSetDevice(c->device));
        DCD_CUDA(cudaStreamSynchronize(c->stream));
        if (c->own_stream) DCD_CUDA(cudaStreamDestroy(c->stream));
        c->stream = (cudaStream_t)cuda_stream;
        c->own_stream = false;
    });
}

extern "C" const char* dcd_last_error(dcd_ctx* c) {
    if (c) return c->err.c_str();
    return g_last_error.c_str();
}

extern "C" int32_t dcd_options_set(dcd_ctx* c, const dcd_options* o) {
    if (!c || !o) return DCD_E_INVALID;
    return guard(c, [&] {
        if (o->min_word_length < 1 || o->max_word_length < o->min_word_length || o->max_word_length >= MAX_KEY_UNITS)
            throw StatusError(DCD_E_INVALID, "word length limits must satisfy 1 <= min <= max <= 47");
        if (o->dict_capacity_log2 && (o->dict_capacity_log2 < 10 || o->dict_capacity_log2 > 28))
            throw StatusError(DCD_E_INVALID, "dict_capacity_log2 must be in [10, 28]");
        c->opt = *o;
        if (!c->opt.dict_capacity_log2) c->opt.dict_capacity_log2 = 23;
    });
}

extern "C" int32_t dcd_counters_get(dcd_ctx* c, dcd_counters* out) {
    if (!c || !out) return DCD_E_INVALID;
    *out = c->counters;
    return DCD_OK;
}

// ---- vocabularies / stop words ---------------------------------------------------------------------
static void upload_vocab(dcd_ctx* c, VocabHost& v) {
    // the vocabulary table is built by the same device insert kernel as the main table
    u32 n = (u32)v.map.size();
    u32 cap_log2 = 10;
    while ((1u << cap_log2) < 2 * n + 16) cap_log2++;
    std::vector<u16> chars;
    std::vector<u32> offs{0}, vals;
    for (auto& kv : v.map) {
        if (kv.first.size() > (size_t)MAX_KEY_UNITS) continue;
        chars.insert(chars.end(), (const u16*)kv.first.data(), (const u16*)kv.first.data() + kv.first.size());
        offs.push_back((u32)chars.size());
        vals.push_back((u32)kv.second);
    }
    n = (u32)vals.size();
    v.dev.create(c->pool, cap_log2, n + 16, (u32)chars.size() + 4 * n + 64, c->stream);
    if (n) {
        DevBuf<u16> dch(c->pool, chars.size());
        DevBuf<u32> dof(c->pool, offs.size()), dv(c->pool, n);
        DCD_CUDA(cudaMemcpyAsync(dch.p, chars.data(), chars.size() * 2, cudaMemcpyHostToDevice, c->stream));
        DCD_CUDA(cudaMemcpyAsync(dof.p, offs.data(), offs.size() * 4, cudaMemcpyHostToDevice, c->stream));
        DCD_CUDA(cudaMemcpyAsync(dv.p, vals.data(), n * 4, cudaMemcpyHostToDevice, c->stream));
        launch_insert_strings(v.dev.view, dch.p, dof.p, n, 0, dv.p, nullptr, c->stream);
        DCD_CUDA(cudaStreamSynchronize(c->stream));
        DCD_CUDA(cudaGetLastError());
    }
}

extern "C" int32_t dcd_vocab_load(dcd_ctx* c, const uint8_t* p, size_t n, int32_t stemmer) {
    if (!c || (!p && n)) return DCD_E_INVALID;
    return guard(c, [&] {
        if (c->vocs.size() >= (size_t)MAX_VOCS) throw StatusError(DCD_E_CAPACITY, "at most 4 vocabularies");
        auto v = std::make_unique<VocabHost>();
        v->stemmer = stemmer;
        // Vocab.Load, Dict.cs:71-95
        size_t i = 0;
        bool have_first = false;
        u16s first_ge_a, last;
        while (i < n) {
            u32 len = 0;
            int sh = 0;
            bool okl = false;
            while (i < n) {
                u8 b = p[i++];
                len |= (u32)(b & 0x7F) << sh;
                sh += 7;
                if (!(b & 0x80)) { okl = true; break; }
            }
            if (!okl || i + len + 4 > n) break;   // EndOfStreamException terminates the loop
            u16s key = utf8_to_utf16(p + i, len);
            i += len;
            int32_t val;
            memcpy(&val, p + i, 4);
            i += 4;
            if (!v->map.emplace(key, val).second) throw StatusError(DCD_E_FORMAT, "duplicate key in .voc");
            v->ordered.emplace_back(key, val);
            if (!have_first && !key.empty() && key[0] >= u'a') { first_ge_a = key; have_first = true; }
            last = key;
        }
        if (!have_first || last.empty()) throw StatusError(DCD_E_FORMAT, ".voc holds no key >= 'a'");
        v->begin = first_ge_a[0];
        v->end = last[0];
        upload_vocab(c, *v);
        c->vocs.push_back(std::move(v));
    });
}

extern "C" int32_t dcd_vocab_add_pairs(dcd_ctx* c, const uint16_t* chars, const uint32_t* offs, const int32_t* groups, int32_t n,
                                       uint16_t rb, uint16_t re, int32_t stemmer) {
    if (!c || n < 0) return DCD_E_INVALID;
    return guard(c, [&] {
        if (c->vocs.size() >= (size_t)MAX_VOCS) throw StatusError(DCD_E_CAPACITY, "at most 4 vocabularies");
        auto v = std::make_unique<VocabHost>();
        v->stemmer = stemmer;
        for (int i = 0; i < n; i++) {
            v->map[mk(chars + offs[i], offs[i + 1] - offs[i])] = groups[i];
            v->ordered.emplace_back(mk(chars + offs[i], offs[i + 1] - offs[i]), groups[i]);
        }
        v->begin = rb;
        v->end = re;
        upload_vocab(c, *v);
        c->vocs.push_back(std::move(v));
    });
}

extern "C" int32_t dcd_stopwords_set(dcd_ctx* c, const uint16_t* chars, const uint32_t* offs, int32_t n) {
    if (!c || n < 0) return DCD_E_INVALID;
    c->stop.clear();
    for (int i = 0; i < n; i++) c->stop.insert(mk(chars + offs[i], offs[i + 1] - offs[i]));
    return DCD_OK;
}

// ===================================================================================================
// build session
// ===================================================================================================
extern "C" int32_t dcd_build_begin(dcd_ctx* c, dcd_build** out) {
    if (!c || !out) return DCD_E_INVALID;
    dcd_build* b = new dcd_build();
    b->ctx = c;
    *out = b;
    return DCD_OK;
}
extern "C" void dcd_build_abort(dcd_build* b) { delete b; }

static PageSegment& explicit_segment(dcd_build* b) {
    if (b->segments.empty() || b->segments.back().bulk) {
        PageSegment sg;
        sg.bulk = false;
        b->segments.push_back(sg);
    }
    return b->segments.back();
}

static void record_page(dcd_build* b, const u16s& id, u64 end) {   // Builder.EndPage(id, coord), Build.cs:364-367
    if (b->ndocs == 0) throw StatusError(DCD_E_STATE, "EndPage before AddDoc");
    PageSegment& sg = explicit_segment(b);
    sg.page_doc.push_back((u32)(b->ndocs - 1));
    sg.page_id.push_back(id);
    sg.count++;
    b->explicit_page_end.push_back(end);
}

static void add_extra(dcd_build* b, const u16s& s, u64 coord, int kind) {
    if (s.empty()) return;
    if (s.size() >= (size_t)MAX_KEY_UNITS) throw StatusError(DCD_E_UNSUPPORTED, "word or code longer than 47 UTF-16 units");
    b->extra_chars.insert(b->extra_chars.end(), (const u16*)s.data(), (const u16*)s.data() + s.size());
    b->extra_offs.push_back((u32)b->extra_chars.size());
    b->extra_coord.push_back(coord);
    b->extra_kind.push_back((u8)kind);
}

// seal the host-staged body text into a device batch so that batches stay in coordinate order
static void flush_host_text(dcd_build* b) {
    if (b->host_page_off.empty()) return;
    dcd_ctx_impl* c = b->ctx;
    TextBatch bt;
    bt.enc = DCD_ENC_UTF16;
    bt.n_units = b->host_text.size();
    bt.npages = b->host_page_off.size();
    b->host_page_off.push_back(b->host_text.size());
    bt.own_text = DevBuf<u8>(c->pool, b->host_text.size() * 2 + 64);
    bt.text = bt.own_text.p;
    bt.own_page_off = DevBuf<u64>(c->pool, b->host_page_off.size());
    bt.page_off = bt.own_page_off.p;
    bt.page_delta = DevBuf<i64>(c->pool, bt.npages);
    DCD_CUDA(cudaMemcpyAsync(bt.own_text.p, b->host_text.data(), b->host_text.size() * 2, cudaMemcpyHostToDevice, c->stream));
    DCD_CUDA(cudaMemcpyAsync(bt.own_page_off.p, b->host_page_off.data(), b->host_page_off.size() * 8, cudaMemcpyHostToDevice, c->stream));
    DCD_CUDA(cudaMemcpyAsync(bt.page_delta.p, b->host_page_delta.data(), bt.npages * 8, cudaMemcpyHostToDevice, c->stream));
    DCD_CUDA(cudaStreamSynchronize(c->stream));
    b->batches.push_back(std::move(bt));
    b->host_text.clear();
    b->host_page_off.clear();
    b->host_page_delta.clear();
}

extern "C" int32_t dcd_build_add_doc(dcd_build* b, const uint16_t* source, int32_t ns, const uint16_t* name, int32_t nn) {
    if (!b || ns < 0 || nn < 0) return DCD_E_INVALID;
    return guard(b->ctx, [&] {
        // Builder.AddDoc, Build.cs:348-356: stored as sourceid + ':' + name
        b->doc_names.resize(b->ndocs + 1);
        b->doc_names[b->ndocs] = mk(source, ns) + u":" + mk(name, nn);
        b->ndocs++;
    });
}

// StringReader.ReadLine
static bool read_line(const u16s& t, size_t& pos, u16s& line) {
    if (pos >= t.size()) return false;
    size_t i = pos;
    while (i < t.size() && t[i] != u'\n' && t[i] != u'\r') i++;
    line = t.substr(pos, i - pos);
    if (i < t.size()) {
        if (t[i] == u'\r' && i + 1 < t.size() && t[i + 1] == u'\n') i++;
        i++;
    }
    pos = i;
    return true;
}

extern "C" int32_t dcd_build_add_page(dcd_build* b, const uint16_t* id_p, int32_t nid, const uint16_t* text, size_t n) {
    if (!b || nid < 0 || (!text && n)) return DCD_E_INVALID;
    return guard(b->ctx, [&] {
        dcd_ctx_impl* c = b->ctx;
        if (n == 0) return;                                        // Build.cs:483
        const u16s id = mk(id_p, nid);
        if (id == u"0") {
            // header page, Build.cs:485-524 (host side: a few hundred bytes per document)
            const u16s page = mk(text, n);
            size_t pos = 0;
            u16s line;
            // a refused word (longer than the device string table holds) must not leave half a page behind: the builder's state
            // is restored before the error goes out (the reference has no length limit on header-field words, Build.cs:505-511;
            // INTEGRATION.md lists the deviation)
            struct Undo {
                dcd_build* b;
                size_t nchars, noffs, ncoord, nkind;
                u64 coord, max_coord;
                bool any, armed = true;
                ~Undo() {
                    if (!armed) return;
                    b->extra_chars.resize(nchars); b->extra_offs.resize(noffs); b->extra_coord.resize(ncoord); b->extra_kind.resize(nkind);
                    b->coord = coord; b->max_coord = max_coord; b->any_posting = any;
                }
            } undo{b, b->extra_chars.size(), b->extra_offs.size(), b->extra_coord.size(), b->extra_kind.size(), b->coord, b->max_coord,
                   b->any_posting};
            while (read_line(page, pos, line)) {
                line = c->lower(line);
                std::vector<u16s> fields;
                size_t st = 0;
                for (size_t i = 0; i <= line.size(); i++)
                    if (i == line.size() || line[i] == u'=') { fields.push_back(line.substr(st, i - st)); st = i + 1; }
                if ((int)fields[0].size() >= c->opt.min_word_length) {
                    if (fields.size() < 2) { undo.armed = false; return; }   // exception path of Build.cs:537: page abandoned without EndPage
                    const u16s& v = fields[1];
                    int dc = (int)fields[0].size() + 1;
                    auto bw = [&](char16_t ch) { return c->is_w(ch) || ch == 0x200C || ch == 0x200D; };   // Regex \b word chars
                    size_t i = 0;
                    while (i < v.size()) {
                        size_t j = i + 1;
                        const bool w = bw(v[i]);
                        while (j < v.size() && bw(v[j]) == w) j++;
                        if (c->is_w(v[i])) {
                            const u64 fc = (u64)(uint32_t)(b->coord + (u64)(dc - 1));   // (uint) cast of Build.cs:509
                            add_extra(b, u"&" + fields[0], fc, 1);
                            b->max_coord = fc;
                            b->any_posting = true;
                            const u16s piece = c->lower(v.substr(i, j - i));
                            add_extra(b, piece, b->coord + (u64)dc, 0);
                            if (!c->word_codes(piece).empty()) b->max_coord = b->coord + (u64)dc;
                        }
                        dc += (int)(j - i);
                        i = j;
                    }
                }
                b->coord += (u64)line.size() + 1;
            }
            record_page(b, id, b->coord);
            undo.armed = false;
            return;
        }
        // body page: staged for the device tokenizer
        if (b->ndocs == 0) throw StatusError(DCD_E_STATE, "page before AddDoc");
        b->host_page_off.push_back(b->host_text.size());
        b->host_page_delta.push_back((i64)b->coord - (i64)b->host_text.size());
        b->host_text.insert(b->host_text.end(), text, text + n);
        b->coord += n;
        b->any_posting = true;
        record_page(b, id, b->coord);
    });
}

extern "C" int32_t dcd_build_add_word(dcd_build* b, const uint16_t* w, int32_t n, uint64_t coord) {
    if (!b || n < 0) return DCD_E_INVALID;
    return guard(b->ctx, [&] {
        const u16s word = mk(w, n);
        add_extra(b, word, coord, 0);
        if (!b->ctx->word_codes(word).empty()) { b->max_coord = coord; b->any_posting = true; }
    });
}
extern "C" int32_t dcd_build_add_posting(dcd_build* b, const uint16_t* code, int32_t n, uint64_t coord) {
    if (!b || n <= 0) return DCD_E_INVALID;
    return guard(b->ctx, [&] {
        add_extra(b, mk(code, n), coord, 1);
        b->max_coord = coord;
        b->any_posting = true;
    });
}
extern "C" int32_t dcd_build_end_page(dcd_build* b, const uint16_t* id, int32_t nid, int32_t use_coord, uint64_t coord) {
    if (!b || nid < 0) return DCD_E_INVALID;
    return guard(b->ctx, [&] { record_page(b, mk(id, nid), use_coord ? coord : b->max_coord); });
}

extern "C" int32_t dcd_build_add_pages_bulk(dcd_build* b, const void* text, int32_t enc, const uint64_t* page_offs, int64_t npages,
                                            const uint64_t* doc_first_page, int64_t ndocs, const uint16_t* source, int32_t ns,
                                            int32_t on_device) {
    if (!b || npages < 0 || ndocs < 0 || (enc != DCD_ENC_LATIN1 && enc != DCD_ENC_UTF16 && enc != DCD_ENC_UTF8)) return DCD_E_INVALID;
    const auto t_begin = std::chrono::steady_clock::now();
    struct AddTimer {
        dcd_build* b; std::chrono::steady_clock::time_point t0;
        ~AddTimer() { b->add_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
    } add_timer{b, t_begin};
    return guard(b->ctx, [&] {
        dcd_ctx_impl* c = b->ctx;
        cudaStream_t s = c->stream;
        if (npages == 0) return;
        flush_host_text(b);
        TextBatch bt;
        bt.enc = enc;
        bt.npages = (u64)npages;
        const int usz = enc == DCD_ENC_UTF16 ? 2 : 1;
        auto seg_dfp = std::make_shared<DevBuf<u64>>(c->pool, (size_t)ndocs + 1);
        u64 first_off = 0, last_off = 0;
        if (on_device) {
            if ((uintptr_t)text & 15) throw StatusError(DCD_E_INVALID, "device text pointer must be 16-byte aligned");
            bt.text = text;
            bt.page_off = reinterpret_cast<const u64*>(page_offs);
            DCD_CUDA(cudaMemcpyAsync(&first_off, page_offs, 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaMemcpyAsync(&last_off, page_offs + npages, 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaMemcpyAsync(seg_dfp->p, doc_first_page, ((size_t)ndocs + 1) * 8, cudaMemcpyDeviceToDevice, s));
            DCD_CUDA(cudaStreamSynchronize(s));
        } else {
            first_off = page_offs[0];
            last_off = page_offs[npages];
            bt.own_text = DevBuf<u8>(c->pool, (size_t)last_off * usz + 64);
            bt.text = bt.own_text.p;
            bt.own_page_off = DevBuf<u64>(c->pool, (size_t)npages + 1);
            bt.page_off = bt.own_page_off.p;
            DCD_CUDA(cudaMemcpyAsync(bt.own_page_off.p, page_offs, ((size_t)npages + 1) * 8, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(seg_dfp->p, doc_first_page, ((size_t)ndocs + 1) * 8, cudaMemcpyHostToDevice, s));
            // the text goes over the copy stream in chunks, one event each: the page-coordinate kernels below and the
            // tokenizer of dcd_build_finish (chunk by chunk) run while later chunks are still in flight
            {
                cudaEvent_t e0;
                DCD_CUDA(cudaEventCreateWithFlags(&e0, cudaEventDisableTiming));
                DCD_CUDA(cudaEventRecord(e0, s));   // the pool hands out memory last used on s
                DCD_CUDA(cudaStreamWaitEvent(c->copy_stream, e0, 0));
                DCD_CUDA(cudaEventDestroy(e0));
                bt.h2d = std::make_shared<TextBatch::H2DChunks>();
                size_t chunk = size_t(96) << 20;
                if (const char* e = getenv("DCD_H2D_CHUNK_KB")) chunk = std::max<size_t>(1, (size_t)atoll(e)) << 10;   // tests: many small chunks
                const size_t total = (size_t)last_off * usz;
                for (size_t at = 0; at < total; at += chunk) {
                    const size_t nb = std::min(chunk, total - at);
                    DCD_CUDA(cudaMemcpyAsync(bt.own_text.p + at, (const u8*)text + at, nb, cudaMemcpyHostToDevice, c->copy_stream));
                    cudaEvent_t ev;
                    DCD_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
                    DCD_CUDA(cudaEventRecord(ev, c->copy_stream));
                    bt.h2d->ev.emplace_back((u64)((at + nb) / usz), ev);
                }
            }
        }
        if (first_off != 0) throw StatusError(DCD_E_INVALID, "page_offs[0] must be 0");
        if (enc == DCD_ENC_UTF8) {
            // UTF-8 pages (page_offs are BYTE offsets): bring the text to the reference's UTF-16 unit axis (utf8.cuh).  Pure ASCII
            // is already DCD_ENC_LATIN1 with the same offsets; otherwise transcode on the device and convert the page offsets.
            if (bt.h2d && !bt.h2d->ev.empty()) DCD_CUDA(cudaStreamWaitEvent(s, bt.h2d->ev.back().second, 0));
            const u64 nbytes = last_off, nblocks = (nbytes + U8_BLOCK - 1) / U8_BLOCK;
            DevBuf<u64> blk_units(c->pool, nblocks + 1), blk_incl(c->pool, nblocks + 1), lb8(c->pool, nblocks / 2048 + 2);
            DevBuf<u32> fl(c->pool, 16);
            DCD_CUDA(cudaMemsetAsync(fl.p, 0, 64, s));
            launch_utf8_scan((const u8*)bt.text, nbytes, blk_units.p, fl.p, c->nsm, s);
            u32 hflags = 0;
            DCD_CUDA(cudaMemcpyAsync(&hflags, fl.p, 4, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            bt.enc = DCD_ENC_LATIN1;
            if (hflags & 1u) {
                const bool wide = (hflags & 2u) != 0;
                DCD_CUDA(cudaMemsetAsync(lb8.p, 0, lb8.bytes(), s));
                DCD_CUDA(cudaMemsetAsync(fl.p, 0, 64, s));
                launch_inclusive_scan_u64(blk_units.p, blk_incl.p + 1, nblocks, 0, lb8.p, fl.p, s);   // blk_incl[k + 1] = units of blocks <= k
                DCD_CUDA(cudaMemsetAsync(blk_incl.p, 0, 8, s));
                u64 total_units = 0;
                DCD_CUDA(cudaMemcpyAsync(&total_units, blk_incl.p + nblocks, 8, cudaMemcpyDeviceToHost, s));
                DCD_CUDA(cudaStreamSynchronize(s));
                DevBuf<u8> out(c->pool, (size_t)total_units * (wide ? 2 : 1) + 64);
                DevBuf<u64> unit_off(c->pool, (size_t)npages + 1);
                launch_utf8_transcode((const u8*)bt.text, nbytes, blk_incl.p, out.p, wide ? 1 : 0, c->nsm, s);
                launch_utf8_page_units((const u8*)bt.text, nbytes, blk_incl.p, total_units, bt.page_off, (u64)npages + 1, unit_off.p, s);
                DCD_CUDA(cudaStreamSynchronize(s));   // the scratch arrays and the UTF-8 copy are released below
                DCD_CUDA(cudaGetLastError());
                bt.h2d.reset();
                bt.own_text = std::move(out);
                bt.text = bt.own_text.p;
                bt.own_page_off = std::move(unit_off);
                bt.page_off = bt.own_page_off.p;
                bt.enc = wide ? DCD_ENC_UTF16 : DCD_ENC_LATIN1;
                last_off = total_units;
            }
        }
        bt.n_units = last_off;
        if (b->coord + last_off > COORD_MASK) throw StatusError(DCD_E_CAPACITY, "coordinate exceeds 2^40");
        // coordinates: adv -> inclusive scan (+ running coord) -> delta
        DevBuf<u64> adv(c->pool, bt.npages);
        bt.coord_end = DevBuf<u64>(c->pool, bt.npages);
        bt.page_delta = DevBuf<i64>(c->pool, bt.npages);
        DevBuf<u64> lb(c->pool, bt.npages / 2048 + 2);
        DevBuf<u32> ticket(c->pool, 16);
        DevBuf<u64> d_cnt(c->pool, 2);
        DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
        DCD_CUDA(cudaMemsetAsync(d_cnt.p, 0, 16, s));
        launch_page_adv(bt.page_off, bt.npages, adv.p, s);
        launch_inclusive_scan_u64(adv.p, bt.coord_end.p, bt.npages, b->coord, lb.p, ticket.p, s);
        launch_page_delta(bt.page_off, adv.p, bt.coord_end.p, nullptr, bt.npages, bt.page_delta.p, s);
        const u64 kept = count_nonzero_u64(adv.p, bt.npages, d_cnt.p, s);
        DCD_CUDA(cudaGetLastError());
        PageSegment sg;
        sg.bulk = true;
        sg.count = kept;
        sg.identity = kept == bt.npages;
        sg.first_doc = b->ndocs;
        sg.ndocs = (u64)ndocs;
        sg.doc_number_base = b->bulk_docs_named;
        sg.source = mk(source, ns);
        sg.npages_orig = bt.npages;
        sg.d_doc_first_page = seg_dfp;
        b->bulk_docs_named += (u64)ndocs;
        b->ndocs += (u64)ndocs;
        b->segments.push_back(sg);   // bulk documents are named implicitly: doc_names stays sparse
        bt.segment = (int)b->segments.size() - 1;
        bt.uniform = true;
        bt.delta0 = (i64)b->coord;
        b->coord += last_off;
        b->any_posting = true;
        b->batches.push_back(std::move(bt));
    });
}

extern "C" int32_t dcd_build_set_origin(dcd_build* b, uint64_t coord_base, uint64_t doc_number_base) {
    if (!b) return DCD_E_INVALID;
    return guard(b->ctx, [&] {
        if (b->any_posting || b->coord != b->coord_origin || b->ndocs)
            throw StatusError(DCD_E_STATE, "dcd_build_set_origin must precede the first document");
        if (coord_base > COORD_MASK) throw StatusError(DCD_E_CAPACITY, "coordinate exceeds 2^40");
        b->coord = b->coord_origin = coord_base;
        b->bulk_docs_named = doc_number_base;
    });
}

extern "C" int32_t dcd_build_finish(dcd_build* b) {
    if (!b) return DCD_E_INVALID;
    dcd_ctx_impl* c = b->ctx;
    int32_t st = guard(c, [&] {
        flush_host_text(b);
        engine_build_finish(b);
    });
    delete b;
    return st;
}

// ===================================================================================================
// sharded build (shard.h)
// ===================================================================================================

extern "C" int32_t dcd_comm_unique_id(uint8_t* out128) {
    if (!out128) return DCD_E_INVALID;
    return guard(nullptr, [&] { nccl_unique_id(out128); });
}
extern "C" int32_t dcd_comm_init_nccl(dcd_ctx* c, const uint8_t* id128, int32_t rank, int32_t nranks, dcd_comm** out) {
    if (!c || !id128 || !out) return DCD_E_INVALID;
    return guard(c, [&] {
        dcd_comm* h = new dcd_comm;
        try { h->impl = make_nccl_comm(c->device, id128, rank, nranks); } catch (...) { delete h; throw; }
        *out = h;
    });
}
extern "C" int32_t dcd_comm_init_local(int32_t nranks, dcd_comm** out) {
    if (!out || nranks < 1 || nranks > 16) return DCD_E_INVALID;
    return guard(nullptr, [&] {
        std::vector<ShardComm*> v(nranks);
        make_local_comms(nranks, v.data());
        for (int r = 0; r < nranks; r++) { out[r] = new dcd_comm; out[r]->impl = v[r]; }
    });
}
extern "C" void dcd_comm_destroy(dcd_comm* comm) {
    if (!comm) return;
    delete comm->impl;
    delete comm;
}
extern "C" int32_t dcd_build_finish_sharded(dcd_build* b, dcd_comm* comm) {
    if (!b || !comm || !comm->impl) return DCD_E_INVALID;
    dcd_ctx_impl* c = b->ctx;
    int32_t st = guard(c, [&] {
        try {
            flush_host_text(b);
            engine_build_finish(b, comm->impl);
        } catch (...) {
            comm->impl->abort();
            throw;
        }
    });
    delete b;
    return st;
}
static thread_local std::vector<uint8_t> t_merged_bytes;
extern "C" int32_t dcd_merged_serialized_size(dcd_ctx* c, uint64_t* ib) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        engine_serialize_merged(c, &t_merged_bytes);
        if (ib) *ib = t_merged_bytes.size();
    });
}
extern "C" int32_t dcd_merged_serialize(dcd_ctx* c, uint8_t* io, uint64_t icap) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        engine_serialize_merged(c, &t_merged_bytes);
        if (t_merged_bytes.size() > icap) throw StatusError(DCD_E_INVALID, "output buffer too small");
        if (io && !t_merged_bytes.empty()) memcpy(io, t_merged_bytes.data(), t_merged_bytes.size());
    });
}

// ===================================================================================================
// persistence
// ===================================================================================================
static thread_local std::vector<uint8_t> t_index_bytes, t_list_bytes;

extern "C" int32_t dcd_index_serialized_size(dcd_ctx* c, uint64_t* ib, uint64_t* lb) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        engine_serialize(c, &t_index_bytes, &t_list_bytes);
        if (ib) *ib = t_index_bytes.size();
        if (lb) *lb = t_list_bytes.size();
    });
}
extern "C" int32_t dcd_index_serialize(dcd_ctx* c, uint8_t* io, uint64_t icap, uint8_t* lo, uint64_t lcap) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        if (t_index_bytes.empty()) engine_serialize(c, &t_index_bytes, &t_list_bytes);
        if (icap < t_index_bytes.size() || lcap < t_list_bytes.size()) throw StatusError(DCD_E_INVALID, "output buffer too small");
        if (io) memcpy(io, t_index_bytes.data(), t_index_bytes.size());
        if (lo) memcpy(lo, t_list_bytes.data(), t_list_bytes.size());
        t_index_bytes.clear();
        t_index_bytes.shrink_to_fit();
        t_list_bytes.clear();
    });
}
static void write_file(const std::string& path, const std::vector<uint8_t>& v) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw StatusError(DCD_E_IO, "cannot create " + path);
    size_t w = v.empty() ? 0 : fwrite(v.data(), 1, v.size(), f);
    fclose(f);
    if (w != v.size()) throw StatusError(DCD_E_IO, "short write to " + path);
}
static std::vector<uint8_t> read_file(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw StatusError(DCD_E_IO, "cannot open " + path);
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> v((size_t)n);
    size_t r = n ? fread(v.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    if (r != (size_t)n) throw StatusError(DCD_E_IO, "short read from " + path);
    return v;
}
extern "C" int32_t dcd_index_save(dcd_ctx* c, const char* dir) {
    if (!c || !dir) return DCD_E_INVALID;
    return guard(c, [&] {
        std::vector<uint8_t> ib, lb;
        engine_serialize(c, &ib, &lb);
        std::string d(dir);
        if (!d.empty() && d.back() != '/') d.push_back('/');
        write_file(d + ".index", ib);          // Index.cs:488, Build.cs:420-422
        write_file(d + ".index.list", lb);     // Index.cs:729-732
    });
}
extern "C" int32_t dcd_index_load_bytes(dcd_ctx* c, const uint8_t* ib, uint64_t ni, const uint8_t* lb, uint64_t nl) {
    if (!c || (!ib && ni) || (!lb && nl)) return DCD_E_INVALID;
    return guard(c, [&] { engine_load_bytes(c, ib, ni, lb, nl); });
}
extern "C" int32_t dcd_index_load(dcd_ctx* c, const char* dir) {
    if (!c || !dir) return DCD_E_INVALID;
    return guard(c, [&] {
        std::string d(dir);
        if (!d.empty() && d.back() != '/') d.push_back('/');
        auto ib = read_file(d + ".index");
        auto lb = read_file(d + ".index.list");
        engine_load_bytes(c, ib.data(), ib.size(), lb.data(), lb.size());
    });
}
extern "C" int32_t dcd_index_sequence(dcd_ctx* c, const uint16_t* code, int32_t n, const uint16_t** data, int32_t* count) {
    if (!c || !data || !count || n < 0) return DCD_E_INVALID;
    return guard(c, [&] {
        if (!c->ix.valid) throw StatusError(DCD_E_STATE, "Index is not built");
        int id = c->find_code(mk(code, n));
        if (id < 0) throw StatusError(DCD_E_INVALID, "KeyNotFound");
        u64 a = c->ix.h_code_off[id], e = c->ix.h_code_off[id + 1];
        c->ix.seq_scratch.resize(e - a);
        if (e > a) DCD_CUDA(cudaMemcpy(c->ix.seq_scratch.data(), c->ix.d_arena.p + a, (e - a) * 2, cudaMemcpyDeviceToHost));
        *data = c->ix.seq_scratch.data();
        *count = (int32_t)(e - a);
    });
}
extern "C" int32_t dcd_index_download_raw(dcd_ctx* c, uint16_t* arena, uint64_t arena_cap, uint64_t* code_offs, uint64_t* page_ends) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        if (!c->ix.valid) throw StatusError(DCD_E_STATE, "Index is not built");
        if (arena) {
            if (arena_cap < c->ix.arena_units) throw StatusError(DCD_E_INVALID, "arena buffer too small");
            DCD_CUDA(cudaMemcpyAsync(arena, c->ix.d_arena.p, c->ix.arena_units * 2, cudaMemcpyDeviceToHost, c->stream));
        }
        if (code_offs) memcpy(code_offs, c->ix.h_code_off.data(), ((size_t)c->ix.ncodes + 1) * 8);
        if (page_ends && c->ix.npages)
            DCD_CUDA(cudaMemcpyAsync(page_ends, c->ix.d_page_end.p, c->ix.npages * 8, cudaMemcpyDeviceToHost, c->stream));
        DCD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

extern "C" int32_t dcd_merged_download_raw(dcd_ctx* c, uint16_t* arena, uint64_t arena_cap, uint64_t* code_offs, uint64_t offs_cap) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        MergedIndex& mg = c->merged;
        if (!mg.valid) throw StatusError(DCD_E_STATE, "no merged index: call dcd_build_finish_sharded first");
        const size_t n_own = mg.code_end - mg.code_begin;
        if (arena && mg.units) {
            if (arena_cap < mg.units) throw StatusError(DCD_E_INVALID, "arena buffer too small");
            DCD_CUDA(cudaMemcpyAsync(arena, mg.d_arena.p, mg.units * 2, cudaMemcpyDeviceToHost, c->stream));
        }
        if (code_offs && mg.d_off.p) {
            if (offs_cap < n_own + 1) throw StatusError(DCD_E_INVALID, "offset buffer too small");
            DCD_CUDA(cudaMemcpyAsync(code_offs, mg.d_off.p, (n_own + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
        }
        DCD_CUDA(cudaStreamSynchronize(c->stream));
    });
}

extern "C" int32_t dcd_word_codes(dcd_ctx* c, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_units,
                                  int32_t* ncodes) {
    if (!c || n < 0) return DCD_E_INVALID;
    return guard(c, [&] {
        auto codes = c->word_codes(mk(word, n));
        int32_t at = 0;
        for (size_t i = 0; i < codes.size(); i++) {
            if (i) { if (at < cap && out) out[at] = 0; at++; }
            for (char16_t ch : codes[i]) { if (at < cap && out) out[at] = (u16)ch; at++; }
        }
        if (out_units) *out_units = at;
        if (ncodes) *ncodes = (int32_t)codes.size();
        if (at > cap && out) throw StatusError(DCD_E_INVALID, "output buffer too small");
    });
}

extern "C" int32_t dcd_index_page_info(dcd_ctx* c, int64_t ordinal, const uint16_t** id, int32_t* nid, int64_t* doc, uint64_t* start,
                                       uint64_t* end) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        u64 d, s0, e0;
        engine_page_info(c, (u64)ordinal, &c->ix.tmp_id, &d, &s0, &e0);
        if (id) *id = (const uint16_t*)c->ix.tmp_id.data();
        if (nid) *nid = (int32_t)c->ix.tmp_id.size();
        if (doc) *doc = (int64_t)d;
        if (start) *start = s0;
        if (end) *end = e0;
    });
}
extern "C" int32_t dcd_index_doc_name(dcd_ctx* c, int64_t doc, const uint16_t** name, int32_t* n) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] {
        c->ix.tmp_name = engine_doc_name(c, (u64)doc);
        if (name) *name = (const uint16_t*)c->ix.tmp_name.data();
        if (n) *n = (int32_t)c->ix.tmp_name.size();
    });
}

extern "C" int32_t dcd_memcpy_d2h(dcd_ctx* c, void* dst, const void* src, uint64_t n) {
    if (!c) return DCD_E_INVALID;
    return guard(c, [&] { DCD_CUDA(cudaMemcpy(dst, src, n, cudaMemcpyDeviceToHost)); });
}

// Host evaluation of the stemmer source shared with the K4 kernel (stem_en.cuh / stem_ru.cuh are host+device).
// Lets the CPU-only test tier compare the product's stemmer with the oracle without a GPU.  No device work.
extern "C" int32_t dcd_stem_host(int32_t stemmer, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_n) {
    if (n < 0 || n > MAX_KEY_UNITS || !out || cap < n || !out_n) return DCD_E_INVALID;
    for (int i = 0; i < n; i++) out[i] = word[i];
    if (stemmer == DCD_STEM_EN) *out_n = en::stem(out, n);
    else if (stemmer == DCD_STEM_RU) *out_n = ru::stem(out, n);
    else *out_n = n;
    return DCD_OK;
}
