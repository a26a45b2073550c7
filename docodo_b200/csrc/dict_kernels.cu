// Dictionary-side queries over the resident key table: every key of the index is scored by one thread,
// the scores are sorted on the device (score in the high 40 bits, key ordinal in the low 24: ties stay in key order, which is
// what the reference's stable OrderBy / Take over the ordinal SortedList gives) and only the first n ordinals come back.
//   GetLikeWords    Search.cs:160-167   '_' -> ".*", unanchored match over keys that start with a letter, first 100
//   GetCloseWords   Search.cs:169-174   10 keys with the smallest Levenshtein distance (Utils.Levenshtein, Index.cs:46-89)
//   GetSuggessions  Search.cs:176-188   keys with the prefix, longer than it, first char >= 'A', by -Value.Count, first n
//   CalcHistogram   Index.cs:284-307    1000 keys with the longest sequences ("#" keys shown as their vocabulary group)
#include <algorithm>
#include <set>

#include "engine.h"
#include "sort.cuh"

namespace dcd {

constexpr u64 SCORE_NONE = ~0ull;
constexpr int KEY_BITS = 24;                       // ordinals < 2^24 (engine.cu refuses more codes)
constexpr int MAX_QUERY_UNITS = 64;                // longest word the Levenshtein kernel takes

DCD_D u64 score_key(u64 score, u32 k) { return (score << KEY_BITS) | k; }

// pattern: the literal segments between the '_' of the request word, packed; a key matches when they occur in order
__global__ void k_like(const u16* __restrict__ chars, const u32* __restrict__ off, u32 ncodes, const u8* __restrict__ cls,
                       const u16* __restrict__ seg_chars, const u32* __restrict__ seg_off, int nseg, u64* __restrict__ out) {
    const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ncodes) return;
    const u16* s = chars + off[k];
    const int n = (int)(off[k + 1] - off[k]);
    bool ok = n > 0 && (cls[s[0]] & CLS_L);
    int pos = 0;
    for (int g = 0; ok && g < nseg; g++) {
        const u16* p = seg_chars + seg_off[g];
        const int m = (int)(seg_off[g + 1] - seg_off[g]);
        int f = -1;
        for (int i = pos; i + m <= n && f < 0; i++) {
            bool eq = true;
            for (int j = 0; j < m && eq; j++) eq = s[i + j] == p[j];
            if (eq) f = i;
        }
        if (f < 0) ok = false; else pos = f + m;
    }
    out[k] = ok ? score_key(0, k) : SCORE_NONE;
}

__global__ void k_levenshtein(const u16* __restrict__ chars, const u32* __restrict__ off, u32 ncodes, const u16* __restrict__ word, int m,
                              u64* __restrict__ out) {
    const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ncodes) return;
    const u16* s = chars + off[k];
    const int n = (int)(off[k + 1] - off[k]);
    int row[MAX_QUERY_UNITS + 1];   // d[i - 1][0..m] -> d[i][0..m]
    u16 t[MAX_QUERY_UNITS];
    for (int j = 0; j < m; j++) t[j] = word[j];
    int dist;
    if (n == 0) dist = m;
    else if (m == 0) dist = n;
    else {
        for (int j = 0; j <= m; j++) row[j] = j;
        for (int i = 1; i <= n; i++) {
            int diag = row[0];
            row[0] = i;
            const u16 si = s[i - 1];
            for (int j = 1; j <= m; j++) {
                const int up = row[j];
                const int v = min(min(up + 1, row[j - 1] + 1), diag + (t[j - 1] == si ? 0 : 1));
                diag = up;
                row[j] = v;
            }
        }
        dist = row[m];
    }
    out[k] = score_key((u64)dist, k);
}

// count_score: (2^40 - 1) - number of ushorts of the key's sequence, so that ascending scores = descending Value.Count
__global__ void k_prefix(const u16* __restrict__ chars, const u32* __restrict__ off, const u64* __restrict__ code_off, u32 ncodes,
                         const u16* __restrict__ word, int m, int any, u64* __restrict__ out) {
    const u32 k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ncodes) return;
    const u16* s = chars + off[k];
    const int n = (int)(off[k + 1] - off[k]);
    bool ok = true;
    if (!any) {
        ok = n > m && s[0] >= 'A';
        for (int j = 0; ok && j < m; j++) ok = s[j] == word[j];
    }
    const u64 cnt = min(code_off[k + 1] - code_off[k], (1ull << 40) - 1);
    out[k] = ok ? score_key(((1ull << 40) - 1) - cnt, k) : SCORE_NONE;
}

static void ensure_device_keys(dcd_ctx_impl* c) {
    ResidentIndex& ix = c->ix;
    if (!ix.valid) throw StatusError(DCD_E_STATE, "Index is not built");
    if (ix.d_key_off.n == (size_t)ix.ncodes + 1) return;
    ix.d_key_chars = DevBuf<u16>(c->pool, ix.h_code_chars.size());
    ix.d_key_off = DevBuf<u32>(c->pool, (size_t)ix.ncodes + 1);
    if (ix.h_code_chars.size())
        DCD_CUDA(cudaMemcpyAsync(ix.d_key_chars.p, ix.h_code_chars.data(), ix.h_code_chars.size() * 2, cudaMemcpyHostToDevice, c->stream));
    DCD_CUDA(cudaMemcpyAsync(ix.d_key_off.p, ix.h_code_str_off.data(), ((size_t)ix.ncodes + 1) * 4, cudaMemcpyHostToDevice, c->stream));
}

// sort the score keys, bring the first `take` ordinals (of keys that matched) back
static std::vector<u32> best_ordinals(dcd_ctx_impl* c, DevBuf<u64>& keys, u32 ncodes, size_t take) {
    cudaStream_t s = c->stream;
    DevBuf<u64> other(c->pool, ncodes), hist(c->pool, 8 * 256);
    DevBuf<u8> scratch(c->pool, sort_scratch_bytes(ncodes));
    const SortPlan plan = make_sort_plan(0, 64);
    DCD_CUDA(cudaMemsetAsync(hist.p, 0, hist.bytes(), s));
    launch_histogram(keys.p, ncodes, plan, hist.p, s);
    u64 launches = 0;
    const int where = radix_sort_run(keys.p, other.p, nullptr, nullptr, ncodes, plan, hist.p, nullptr, scratch.p, s, &launches);
    const u64* sorted = where ? other.p : keys.p;
    const size_t n = std::min<size_t>(take, ncodes);
    std::vector<u64> h(n);
    if (n) DCD_CUDA(cudaMemcpyAsync(h.data(), sorted, n * 8, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    DCD_CUDA(cudaGetLastError());
    std::vector<u32> out;
    for (u64 v : h) {
        if (v == SCORE_NONE) break;
        out.push_back((u32)(v & ((1u << KEY_BITS) - 1)));
    }
    return out;
}

static unsigned grid_of(u32 n) { return (n + 127) / 128; }

std::vector<u16s> engine_like_words(dcd_ctx_impl* c, const u16s& word) {
    if (word.find(u'_') == u16s::npos || !c->opt.keep_forms) return {word};
    if (word.size() < 2) return {};
    ensure_device_keys(c);
    ResidentIndex& ix = c->ix;
    if (!ix.ncodes) return {};
    std::vector<u16> sc;
    std::vector<u32> so{0};
    size_t st = 0;
    for (size_t i = 0; i <= word.size(); i++)
        if (i == word.size() || word[i] == u'_') {
            if (i > st) { sc.insert(sc.end(), (const u16*)word.data() + st, (const u16*)word.data() + i); so.push_back((u32)sc.size()); }
            st = i + 1;
        }
    cudaStream_t s = c->stream;
    DevBuf<u16> d_sc(c->pool, sc.size() + 1);
    DevBuf<u32> d_so(c->pool, so.size());
    DevBuf<u64> keys(c->pool, ix.ncodes);
    if (!sc.empty()) DCD_CUDA(cudaMemcpyAsync(d_sc.p, sc.data(), sc.size() * 2, cudaMemcpyHostToDevice, s));
    DCD_CUDA(cudaMemcpyAsync(d_so.p, so.data(), so.size() * 4, cudaMemcpyHostToDevice, s));
    k_like<<<grid_of(ix.ncodes), 128, 0, s>>>(ix.d_key_chars.p, ix.d_key_off.p, ix.ncodes, c->d_cls.p, d_sc.p, d_so.p, (int)so.size() - 1, keys.p);
    std::vector<u16s> out;
    for (u32 k : best_ordinals(c, keys, ix.ncodes, 100)) out.push_back(c->code_string(k));   // MAX_LIKE_WORDS
    return out;
}

std::vector<u16s> engine_close_words(dcd_ctx_impl* c, const u16s& word) {
    if (word.size() > (size_t)MAX_QUERY_UNITS) throw StatusError(DCD_E_UNSUPPORTED, "GetCloseWords: words of at most 64 characters");
    ensure_device_keys(c);
    ResidentIndex& ix = c->ix;
    if (!ix.ncodes) return {};
    cudaStream_t s = c->stream;
    DevBuf<u16> d_w(c->pool, word.size() + 1);
    DevBuf<u64> keys(c->pool, ix.ncodes);
    if (!word.empty()) DCD_CUDA(cudaMemcpyAsync(d_w.p, word.data(), word.size() * 2, cudaMemcpyHostToDevice, s));
    k_levenshtein<<<grid_of(ix.ncodes), 128, 0, s>>>(ix.d_key_chars.p, ix.d_key_off.p, ix.ncodes, d_w.p, (int)word.size(), keys.p);
    std::vector<u16s> out;
    for (u32 k : best_ordinals(c, keys, ix.ncodes, 10)) out.push_back(c->code_string(k));
    return out;
}

std::vector<u16s> engine_suggestions(dcd_ctx_impl* c, const u16s& req, int n) {
    if (req.size() < 2 || n <= 0) return {};
    // Regex.Split(req, @"\b").Last(s => s.Length > 0).ToLower()
    auto bw = [&](char16_t ch) { return c->is_w(ch) || ch == 0x200C || ch == 0x200D; };
    size_t st = 0;
    u16s last;
    for (size_t i = 1; i <= req.size(); i++)
        if (i == req.size() || bw(req[i - 1]) != bw(req[i])) { if (i > st) last = req.substr(st, i - st); st = i; }
    const u16s lastword = c->lower(last);
    if (lastword.size() < 2) return {};
    ensure_device_keys(c);
    ResidentIndex& ix = c->ix;
    if (!ix.ncodes) return {};
    cudaStream_t s = c->stream;
    DevBuf<u16> d_w(c->pool, lastword.size() + 1);
    DevBuf<u64> keys(c->pool, ix.ncodes);
    DCD_CUDA(cudaMemcpyAsync(d_w.p, lastword.data(), lastword.size() * 2, cudaMemcpyHostToDevice, s));
    k_prefix<<<grid_of(ix.ncodes), 128, 0, s>>>(ix.d_key_chars.p, ix.d_key_off.p, ix.d_code_off.p, ix.ncodes, d_w.p, (int)lastword.size(), 0, keys.p);
    std::vector<u16s> out;
    for (u32 k : best_ordinals(c, keys, ix.ncodes, (size_t)n)) out.push_back(c->code_string(k).substr(lastword.size()));
    return out;
}

std::vector<std::pair<u16s, u64>> engine_histogram(dcd_ctx_impl* c, int top) {
    ensure_device_keys(c);
    ResidentIndex& ix = c->ix;
    std::vector<std::pair<u16s, u64>> out;
    if (!ix.ncodes || top <= 0) return out;
    cudaStream_t s = c->stream;
    DevBuf<u64> keys(c->pool, ix.ncodes);
    k_prefix<<<grid_of(ix.ncodes), 128, 0, s>>>(ix.d_key_chars.p, ix.d_key_off.p, ix.d_code_off.p, ix.ncodes, nullptr, 0, 1, keys.p);
    std::set<u16s> seen;
    for (u32 k : best_ordinals(c, keys, ix.ncodes, (size_t)top)) {
        const u16s key = c->code_string(k);
        u16s shown = key;
        if (!key.empty() && key[0] == u'#') {   // "(" + GetWordsGroup(code) + ")", Index.cs:270-282: first 20 words of the group
            unsigned long long id = 0;
            bool okhex = key.size() > 1;
            for (size_t i = 1; i < key.size() && okhex; i++) {
                const char16_t ch = key[i];
                const int d = ch >= u'0' && ch <= u'9' ? ch - u'0' : ch >= u'A' && ch <= u'F' ? ch - u'A' + 10 : ch >= u'a' && ch <= u'f' ? ch - u'a' + 10 : -1;
                if (d < 0) okhex = false; else id = id * 16 + (unsigned)d;
            }
            const size_t nV = (size_t)(id >> 24);
            u16s words;
            int taken = 0;
            if (okhex && nV < c->vocs.size())
                for (const auto& kv : c->vocs[nV]->ordered) {
                    if (kv.second != (int32_t)(id & 0xFFFFFF)) continue;
                    if (taken++) words += u",";
                    words += kv.first;
                    if (taken >= 20) break;
                }
            if (!taken) break;   // Aggregate over an empty group throws: the reference's loop ends there (Index.cs:302-305)
            shown = u"(" + words + u")";
        }
        if (!seen.insert(shown).second) break;   // Dictionary.Add of a duplicate display key throws: same
        out.emplace_back(shown, ix.h_code_off[k + 1] - ix.h_code_off[k]);
    }
    return out;
}

}  // namespace dcd
