// Index.Search on the device: host planning + batched K8/K9 execution + result shaping.
//   request rewriting      PrepareSearchRequest, Search.cs:319-363 (hand-written equivalents of its regular expressions)
//   term resolution        SearchWord / SearchField / GetLikeWords, Search.cs:126-167,192-260; SearchSequence.d(), 302-312
//   expression             C# precedence as evaluated by DynamicExpresso (Search.cs:501,529): '*' over '+', parentheses
//   result shaping         PrepareSearchResult / doc ranks / ordering, Search.cs:365-428,552-557,599-603
#include "search.h"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <cstdio>
#include <cstdlib>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <regex>
#include <unordered_set>

#include "encode.cuh"
#include "query_kernels.cuh"
#include "shard.h"

namespace dcd {

namespace {

struct Term {
    u16s word;      // as typed (after rewriting): upper-case means exact, '_' means wildcard
    int kind = 0;   // 0: SearchWord, 1: SearchField(field, word)
    u16s field;
    int nfound_slot = -1;
};

struct Rewriter {
    const dcd_ctx_impl* c;
    std::vector<Term>* terms;

    bool w(char16_t ch) const { return c->is_w(ch); }
    bool bw(char16_t ch) const { return c->is_w(ch) || ch == 0x200C || ch == 0x200D; }   // Regex \b word characters
    bool bnd(const u16s& s, size_t i) const {
        const bool a = i > 0 && bw(s[i - 1]);
        const bool b = i < s.size() && bw(s[i]);
        return a != b;
    }

    // Returns the expression with one private-use char (U+E000 + term index) per variable.
    u16s run(u16s req, u16s* fields, int kind, const u16s& field, bool keepshort) {
        fields->clear();
        {   // Search.cs:325   [^\w(){}=~?|"]|_+  ->  " "
            u16s o;
            for (size_t i = 0; i < req.size();) {
                const char16_t ch = req[i];
                if (ch == u'_') {
                    while (i < req.size() && req[i] == u'_') i++;
                    o.push_back(u' ');
                    continue;
                }
                const bool keep = w(ch) || ch == u'(' || ch == u')' || ch == u'{' || ch == u'}' || ch == u'=' || ch == u'~' ||
                                  ch == u'?' || ch == u'|' || ch == u'"';
                o.push_back(keep ? ch : u' ');
                i++;
            }
            req.swap(o);
        }
        u16s f;
        {   // Search.cs:328-333   {*(\w+)[ ]*=([\w|() ]+)}   -> recursive field expression, removed from the request
            u16s o;
            size_t i = 0;
            while (i < req.size()) {
                size_t j = i;
                while (j < req.size() && req[j] == u'{') j++;
                size_t k = j;
                while (k < req.size() && w(req[k])) k++;
                bool ok = k > j;
                size_t m = k;
                if (ok) {
                    while (m < req.size() && req[m] == u' ') m++;
                    ok = m < req.size() && req[m] == u'=';
                }
                size_t n = m + 1;
                if (ok) {
                    while (n < req.size() && (w(req[n]) || req[n] == u'|' || req[n] == u'(' || req[n] == u')' || req[n] == u' ')) n++;
                    ok = n > m + 1 && n < req.size() && req[n] == u'}';
                }
                if (ok) {
                    u16s dummy;
                    const u16s l = run(req.substr(m + 1, n - m - 1), &dummy, 1, req.substr(j, k - j), true);
                    if (!f.empty()) f.push_back(u'*');
                    f += u"(" + l + u")";
                    i = n + 1;
                } else {
                    o.push_back(req[i++]);
                }
            }
            req.swap(o);
        }
        *fields = f;
        {   // Search.cs:337   {.*} -> ""
            const size_t a = req.find(u'{'), b = req.rfind(u'}');
            if (a != u16s::npos && b != u16s::npos && b > a) req.erase(a, b - a + 1);
        }
        for (auto& ch : req) if (ch == u'?') ch = u'_';   // Search.cs:340
        if (!keepshort) {   // Search.cs:343-344   \b\w{1,2}\b -> " "
            u16s o;
            for (size_t i = 0; i < req.size();) {
                if (w(req[i])) {
                    size_t j = i;
                    while (j < req.size() && w(req[j])) j++;
                    if (j - i <= 2 && bnd(req, i) && bnd(req, j)) o.push_back(u' ');
                    else o.append(req, i, j - i);
                    i = j;
                } else {
                    o.push_back(req[i++]);
                }
            }
            req.swap(o);
        }
        for (const auto& st : c->stop) {   // Search.cs:346   \b<stop>\b -> ""
            if (st.empty()) continue;
            u16s o;
            size_t i = 0;
            while (i < req.size()) {
                if (req.compare(i, st.size(), st) == 0 && bnd(req, i) && bnd(req, i + st.size())) i += st.size();
                else o.push_back(req[i++]);
            }
            req.swap(o);
        }
        {   // Search.cs:349   "(.*)" -> (UPPER)
            const size_t a = req.find(u'"'), b = req.rfind(u'"');
            if (a != u16s::npos && b != u16s::npos && b > a)
                req = req.substr(0, a) + u"(" + c->upper(req.substr(a + 1, b - a - 1)) + u")" + req.substr(b + 1);
        }
        for (auto& ch : req) if (ch == u'|') ch = u'+';   // Search.cs:351
        {   // Search.cs:353   (\b|\))(\s+)(\b|\()  : the blank run becomes '*'
            u16s o;
            size_t i = 0;
            while (i < req.size()) {
                size_t ws0 = 0;
                bool lead = false, ok = false;
                if (req[i] == u' ' && bnd(req, i)) { ok = true; ws0 = i; }
                else if (req[i] == u')' && i + 1 < req.size() && req[i + 1] == u' ') { ok = true; lead = true; ws0 = i + 1; }
                if (ok) {
                    size_t e = ws0;
                    while (e < req.size() && req[e] == u' ') e++;
                    const bool tail_b = e < req.size() && bnd(req, e);
                    const bool tail_p = !tail_b && e < req.size() && req[e] == u'(';
                    if (tail_b || tail_p) {
                        if (lead) o.push_back(u')');
                        o.push_back(u'*');
                        if (tail_p) { o.push_back(u'('); e++; }
                        i = e;
                        continue;
                    }
                }
                o.push_back(req[i++]);
            }
            req.swap(o);
        }
        {   // Search.cs:357   \b(\w+)\b -> variable
            u16s o;
            for (size_t i = 0; i < req.size();) {
                if (w(req[i])) {
                    size_t j = i;
                    while (j < req.size() && w(req[j])) j++;
                    Term t;
                    t.word = req.substr(i, j - i);
                    t.kind = kind;
                    t.field = field;
                    terms->push_back(t);
                    o.push_back((char16_t)(0xE000 + terms->size() - 1));
                    i = j;
                } else {
                    o.push_back(req[i++]);
                }
            }
            req.swap(o);
        }
        return req;
    }
};

enum { PF_VAR = 0, PF_MUL = 1, PF_ADD = 2 };
struct PfTok { int type; int var; };

// recursive descent with C# precedence; returns false on a syntax error
struct ExprParser {
    const u16s& s;
    size_t nterms;
    size_t i = 0;
    std::vector<PfTok>* out;
    bool err = false;
    void ws() { while (i < s.size() && s[i] == u' ') i++; }
    void primary() {
        ws();
        if (i >= s.size()) { err = true; return; }
        const char16_t ch = s[i];
        if (ch == u'(') {
            i++;
            expr();
            ws();
            if (err || i >= s.size() || s[i] != u')') { err = true; return; }
            i++;
            return;
        }
        if (ch >= 0xE000 && (size_t)(ch - 0xE000) < nterms) { out->push_back({PF_VAR, ch - 0xE000}); i++; return; }
        err = true;
    }
    void term() {
        primary();
        while (!err) {
            ws();
            if (i < s.size() && s[i] == u'*') { i++; primary(); if (!err) out->push_back({PF_MUL, 0}); }
            else break;
        }
    }
    void expr() {
        term();
        while (!err) {
            ws();
            if (i < s.size() && s[i] == u'+') { i++; term(); if (!err) out->push_back({PF_ADD, 0}); }
            else break;
        }
    }
    bool parse() {
        expr();
        ws();
        return !err && i == s.size();
    }
};

struct Plan {   // one device program
    std::vector<QInstr> instr;
    std::vector<u32> codes;
    u64 list_units = 0;   // total list area
    u64 tmp_cap = 1;
    u64 hit_cap = 0;
    u64 rmax = 0;         // largest |R| of any list of the program: bound of the gaps inside a run of any operator*
    bool has_mul = false;
    // early termination (QDesc::limit): level 0 / 1 = lists decoded up to the end of page 2 / 32 x (MAX_FOUND_PAGES + 1),
    // 2 = whole lists.  base_instr keeps the planner's layout; instr is re-laid for the level's list capacities.
    std::vector<QInstr> base_instr;
    int level = 2;
    u64 limit = ~0ull;
    u64 work = 0;         // elements the program touches (decoded lists + operator inputs): launch order, longest first
};

// Lay the program's list area out for the given capacity of every LOAD (same arithmetic as Planner): LOAD regions of one half
// (two for an OR of several codes), operator outputs of the summed capacities.
// A cached LOAD (QOP_LOADC) reads its list in place: it gets a region only when copy_cached says so (a halo goes in front).
static void relayout(Plan& pl, const std::vector<u64>& load_cap, const std::vector<u8>* copy_cached = nullptr) {
    std::vector<u64> st;
    u64 units = 0, tmpc = 1, work = 0;
    size_t li = 0;
    for (auto& in : pl.instr) {
        if (in.op == QOP_LOADC) {
            const u64 cap = load_cap[li];
            in.out_off = units;
            if (copy_cached && (*copy_cached)[li]) units += cap;
            li++;
            st.push_back(cap);
        } else if (in.op == QOP_LOAD) {
            const u64 cap = load_cap[li++];
            if (cap > 0xFFFFFFFFull) throw StatusError(DCD_E_CAPACITY, "posting lists of one query exceed 2^32 entries");
            if (in.b > 1) tmpc = std::max<u64>(tmpc, in.cap);
            in.cap = (u32)cap;
            in.out_off = units;
            units += (u64)(in.b > 1 ? 2 : 1) * cap;
            work += cap;
            st.push_back(cap);
        } else if (in.op == QOP_MUL || in.op == QOP_ADD) {
            const u64 b = st.back(); st.pop_back();
            const u64 a = st.back(); st.pop_back();
            in.out_off = units;
            units += a + b;
            // operator*: two merge passes, unless one side is short enough for galloping
            work += in.op == QOP_ADD ? a + b : (std::min(a, b) * 12 <= std::max(a, b) ? std::min(a, b) * 8 : 2 * (a + b));
            tmpc = std::max(tmpc, a + b);
            st.push_back(a + b);
        }
    }
    pl.work = work;
    pl.list_units = units;
    pl.tmp_cap = tmpc;
    pl.hit_cap = st.empty() ? 0 : st.back();
}

struct Planner {
    const dcd_ctx_impl* c;
    int dist;
    Plan* p;
    std::vector<Term>* terms;
    struct Ent { int R; u64 cap; };
    std::vector<Ent> stack;

    u64 code_len(u32 id) const { return c->ix.h_code_off[id + 1] - c->ix.h_code_off[id]; }

    // GetLikeWords, Search.cs:160-167
    std::vector<u16s> like_words(const u16s& word) const {
        if (word.find(u'_') == u16s::npos || !c->opt.keep_forms) return {word};
        if (word.size() < 2) return {};
        std::vector<u16s> segs;
        size_t st = 0;
        for (size_t i = 0; i <= word.size(); i++)
            if (i == word.size() || word[i] == u'_') { if (i > st) segs.push_back(word.substr(st, i - st)); st = i + 1; }
        std::vector<u16s> out;
        const ResidentIndex& ix = c->ix;
        for (u32 k = 0; k < ix.ncodes && out.size() < 100; k++) {
            const u16* s = ix.h_code_chars.data() + ix.h_code_str_off[k];
            const size_t n = ix.h_code_str_off[k + 1] - ix.h_code_str_off[k];
            if (!n || !c->is_L((char16_t)s[0])) continue;
            const u16s key((const char16_t*)s, n);
            size_t pos = 0;
            bool ok = true;
            for (const auto& sg : segs) {
                const size_t f = key.find(sg, pos);
                if (f == u16s::npos) { ok = false; break; }
                pos = f + sg.size();
            }
            if (ok) out.push_back(key);
        }
        return out;
    }

    // SearchWord, Search.cs:192-260: emits one LOAD; returns R of `total` (-1 exact, 0 otherwise)
    int search_word(const u16s& word_in) {
        u16s word = word_in;
        bool exact = c->upper(word) == word && c->opt.keep_forms;
        word = c->lower(word);
        std::vector<u16s> words{word};
        std::vector<u32> ids;
        bool dead = false;
        if (word.find(u'_') != u16s::npos) {
            if (c->opt.keep_forms) { exact = true; words = like_words(word); }
            else dead = true;   // returns new IndexSequence() before bExact is applied
        }
        if (!dead) {
            for (const auto& ww : words) {
                const std::vector<u16s> codes = c->word_codes(ww);
                if (codes.empty()) continue;
                std::vector<u16s> know;
                const u16s* self = nullptr;
                for (const auto& s : codes) {
                    if (c->is_w(s[0])) { if (!self) self = &s; }
                }
                for (const auto& s : codes) {
                    if (c->is_w(s[0])) continue;
                    if (std::find(know.begin(), know.end(), s) == know.end()) know.push_back(s);
                }
                std::vector<u16s> use;
                if (!exact && !know.empty()) use = know;
                else if (self) use.push_back(*self);
                for (const auto& code : use) {
                    const int id = c->find_code(code);
                    if (id >= 0) ids.push_back((u32)id);
                }
            }
        }
        QInstr in{};
        in.op = QOP_LOAD;
        in.a = (u32)p->codes.size();
        in.b = (u32)ids.size();
        u64 cap = 0;
        for (u32 id : ids) { p->codes.push_back(id); cap += code_len(id); }
        in.cap = (u32)std::min<u64>(cap, 0xFFFFFFFFull);
        in.out_off = p->list_units;
        p->list_units += (ids.size() > 1 ? 2 : 1) * cap;
        if (ids.size() > 1) p->tmp_cap = std::max(p->tmp_cap, cap);
        p->instr.push_back(in);
        stack.push_back({dead ? 0 : (exact ? -1 : 0), cap});
        return stack.back().R;
    }

    void binop(int op) {
        Ent b = stack.back(); stack.pop_back();
        Ent a = stack.back(); stack.pop_back();
        const int absR = std::max(std::abs(a.R), std::abs(b.R));   // IndexSequence.cs:221-222 / 288-289
        const int R = (a.R < 0 && b.R < 0) ? -absR : absR;
        if (op == PF_MUL) {
            p->has_mul = true;
            p->rmax = std::max<u64>(p->rmax, absR == 0 ? (1ull << 62) : (u64)absR);   // absR == 0: the whole union is one run
        }
        QInstr in{};
        in.op = op == PF_MUL ? QOP_MUL : QOP_ADD;
        in.a = (u32)absR;
        in.b = R < 0 ? 1u : 0u;
        in.out_off = p->list_units;
        const u64 cap = a.cap + b.cap;
        p->list_units += cap;
        p->tmp_cap = std::max(p->tmp_cap, cap);
        p->instr.push_back(in);
        stack.push_back({R, cap});
    }

    // SearchSequence.d(), Search.cs:302-312
    void variable(int v) {
        Term& t = (*terms)[v];
        if (t.kind == 0) {
            search_word(t.word);
        } else {
            // SearchField, Search.cs:126-155
            const int fid = c->find_code(u"&" + c->lower(t.field));
            if (fid < 0) {
                QInstr in{};
                in.op = QOP_LOAD;
                in.a = (u32)p->codes.size();
                in.b = 0;
                in.out_off = p->list_units;
                p->instr.push_back(in);
                stack.push_back({0, 0});
            } else {
                QInstr in{};
                in.op = QOP_LOAD;
                in.a = (u32)p->codes.size();
                in.b = 1;
                p->codes.push_back((u32)fid);
                const u64 cap = code_len((u32)fid);
                in.cap = (u32)cap;
                in.out_off = p->list_units;
                p->list_units += cap;
                p->instr.push_back(in);
                stack.push_back({-1, cap});          // seq.R = -1
                search_word(c->lower(t.word));
                binop(PF_MUL);
            }
        }
        Ent& top = stack.back();
        top.R = top.R < 0 ? -(int)t.word.size() - 4 : dist + (int)t.word.size();   // Search.cs:308
        QInstr cnt{};
        cnt.op = QOP_COUNT;
        cnt.a = (u32)t.nfound_slot;
        p->instr.push_back(cnt);
    }

    bool build(const std::vector<PfTok>& pf) {
        for (const auto& t : pf) {
            if (t.type == PF_VAR) variable(t.var);
            else {
                if (stack.size() < 2) return false;
                binop(t.type);
            }
            if (stack.size() > 14) return false;
        }
        if (stack.size() != 1) return false;
        p->hit_cap = stack.back().cap;
        return true;
    }
};

struct Prepared {
    int status = 0;
    std::vector<Term> terms;
    std::vector<u16s> filter;
    int main_plan = -1, field_plan = -1;
};

}  // namespace

// k-th string of a packed segment list copied on the device: dst[seg_dst[i] .. +seg_cnt[i]) = src[seg_src[i] ..)
static void copy_segments(DevPool& pool, const u64* src, u64* dst, const std::vector<u64>& ssrc, const std::vector<u64>& sdst,
                          const std::vector<u32>& scnt, cudaStream_t s) {
    if (ssrc.empty()) return;
    DevBuf<u64> d_src(pool, ssrc.size()), d_dst(pool, sdst.size());
    DevBuf<u32> d_cnt(pool, scnt.size());
    DCD_CUDA(cudaMemcpyAsync(d_src.p, ssrc.data(), ssrc.size() * 8, cudaMemcpyHostToDevice, s));
    DCD_CUDA(cudaMemcpyAsync(d_dst.p, sdst.data(), sdst.size() * 8, cudaMemcpyHostToDevice, s));
    DCD_CUDA(cudaMemcpyAsync(d_cnt.p, scnt.data(), scnt.size() * 4, cudaMemcpyHostToDevice, s));
    launch_copy_segments(src, dst, d_src.p, d_dst.p, d_cnt.p, (u32)ssrc.size(), s);
    DCD_CUDA(cudaStreamSynchronize(s));   // the descriptor arrays are released on return
}

// Exact serving from document-range shards (query partitioning): per code, the first coordinate on the LATER
// shards.  One all-gather per resident index, at its first sharded batch.
static void ensure_next_first(dcd_ctx_impl* c, ShardComm* comm, cudaStream_t s) {
    ResidentIndex& ix = c->ix;
    if (ix.d_next_first.n == (size_t)ix.ncodes + 1 && ix.next_first_ranks == comm->nranks) return;
    const int R = comm->nranks;
    const size_t row = (size_t)ix.ncodes + 1;   // + this shard's origin
    DevBuf<u64> mine(c->pool, row), all(c->pool, row * R);
    launch_first_coords(ix.d_arena.p, ix.d_code_off.p, ix.d_code_base.n ? ix.d_code_base.p : nullptr, ix.ncodes, mine.p, s);
    DCD_CUDA(cudaMemcpyAsync(mine.p + ix.ncodes, &ix.coord_origin, 8, cudaMemcpyHostToDevice, s));
    comm->allgather(mine.p, all.p, row * 8, s);
    ix.d_next_first = DevBuf<u64>(c->pool, row);
    launch_next_first(all.p, R, comm->rank, (u32)row, ix.ncodes, ix.d_next_first.p, s);
    ix.shard_origins.assign((size_t)R, 0);
    for (int r = 0; r < R; r++)
        DCD_CUDA(cudaMemcpyAsync(&ix.shard_origins[r], all.p + (size_t)r * row + ix.ncodes, 8, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    DCD_CUDA(cudaGetLastError());
    ix.next_first_ranks = R;
}

// number of coordinates of every code's list (wordInfo.nFound of a single-code term), once per resident index
static void ensure_code_nvals(dcd_ctx_impl* c, cudaStream_t s) {
    ResidentIndex& ix = c->ix;
    if (ix.h_code_nvals.size() == (size_t)ix.ncodes) return;
    DevBuf<u32> d(c->pool, std::max<u32>(ix.ncodes, 1));
    launch_code_nvals(ix.d_arena.p, ix.d_code_off.p, ix.ncodes, d.p, s);
    ix.h_code_nvals.assign(ix.ncodes, 0);
    if (ix.ncodes) DCD_CUDA(cudaMemcpyAsync(ix.h_code_nvals.data(), d.p, (size_t)ix.ncodes * 4, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    DCD_CUDA(cudaGetLastError());
}

void engine_search_batch(dcd_ctx_impl* c, const std::vector<u16s>& reqs, const std::vector<int>& has_opt, const std::vector<int>& dist,
                         dcd_hits* out, PagesSink* sink, ShardComm* comm) {
    const size_t nq = reqs.size();
    out->q.assign(nq, QueryResult());
    if (!c->ix.valid) {
        for (auto& r : out->q) r.status = DCD_E_STATE;   // "Index is not built", Search.cs:445
        return;
    }
    cudaStream_t s = c->stream;
    DevPool& pool = c->pool;
    const int trace = getenv("DCD_SEARCH_TIMING") ? std::max(1, atoi(getenv("DCD_SEARCH_TIMING"))) : 0;   // host phase times on stderr (development aid; 2: longest programs)
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto since = [&](std::chrono::steady_clock::time_point t) { return std::chrono::duration<double, std::milli>(now() - t).count(); };
    auto t_phase = now();
    double ms_plan = 0, ms_wave_prep = 0, ms_kernel_wait = 0, ms_download = 0, ms_shape = 0;
    std::vector<Prepared> prep(nq);
    std::vector<Plan> plans;
    u32 nslots = 0;

    // ---- host planning ------------------------------------------------------------------------------
    // Requests are planned independently on several host threads (the rewriter, GetWordCodes and the key-table look-ups only
    // read the context); term slots and plan numbers are local to the request until the serial merge below.
    std::vector<std::vector<Plan>> local_plans(nq);
    auto plan_query = [&](size_t qi) {
        Prepared& pr = prep[qi];
        std::vector<Plan>& plans = local_plans[qi];
        u32 nslots = 0;
        u16s req = c->lower(reqs[qi]);   // Search.cs:453
        {   // -filter: (Search.cs:456-466)
            const u16s tag = u"-filter:";
            auto fchar = [&](char16_t ch) {
                return c->is_w(ch) || ch == u'*' || ch == u'?' || ch == u'\\' || ch == u'.' || ch == u'(' || ch == u')' || ch == u'+' ||
                       ch == u'{' || ch == u'}' || ch == u'/';
            };
            auto bw = [&](char16_t ch) { return c->is_w(ch) || ch == 0x200C || ch == 0x200D; };
            u16s o;
            size_t i = 0;
            bool first = true;
            while (i < req.size()) {
                const bool nb = !((i > 0 && bw(req[i - 1])) != (i < req.size() && bw(req[i])));   // \B
                if (req.compare(i, tag.size(), tag) == 0 && nb) {
                    size_t j = i + tag.size();
                    std::vector<u16s> caps;
                    while (j < req.size() && fchar(req[j])) {
                        size_t k = j;
                        while (k < req.size() && fchar(req[k])) k++;
                        caps.push_back(req.substr(j, k - j));
                        if (k < req.size() && req[k] == u',') k++;
                        j = k;
                    }
                    if (!caps.empty()) {
                        if (first) { pr.filter = caps; first = false; }
                        o.push_back(u' ');
                        i = j;
                        continue;
                    }
                }
                o.push_back(req[i++]);
            }
            req.swap(o);
        }
        u16s fields;
        Rewriter rw{c, &pr.terms};
        const u16s expr = rw.run(req, &fields, 0, u16s(), false);   // Search.cs:470
        if (pr.terms.size() > 4096) { pr.status = DCD_E_SYNTAX; return; }
        for (auto& t : pr.terms) t.nfound_slot = (int)nslots++;
        const int R = has_opt[qi] ? dist[qi] : 255;                   // Search.cs:472-473
        auto make_plan = [&](const u16s& e) -> int {
            std::vector<PfTok> pf;
            ExprParser ep{e, pr.terms.size()};
            ep.out = &pf;
            if (!ep.parse()) return -2;
            Plan pl;
            Planner pn{c, R, &pl, &pr.terms};
            if (!pn.build(pf)) return -2;
            plans.push_back(std::move(pl));
            return (int)plans.size() - 1;
        };
        if (!expr.empty()) {
            pr.main_plan = make_plan(expr);
            if (pr.main_plan == -2) { pr.status = DCD_E_SYNTAX; return; }
        }
        if (!fields.empty()) {
            pr.field_plan = make_plan(fields);
            if (pr.field_plan == -2) { pr.status = DCD_E_SYNTAX; return; }
        }
    };
    {
        unsigned nth = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
        if (nq < 64) nth = 1;
        if (nth <= 1) {
            for (size_t qi = 0; qi < nq; qi++) plan_query(qi);
        } else {
            std::atomic<size_t> next{0};
            std::vector<std::thread> th;
            std::vector<std::exception_ptr> errs(nth);
            for (unsigned ti = 0; ti < nth; ti++)
                th.emplace_back([&, ti] {
                    try {
                        for (;;) {
                            const size_t q0 = next.fetch_add(32);
                            if (q0 >= nq) break;
                            for (size_t qi = q0; qi < std::min(nq, q0 + 32); qi++) plan_query(qi);
                        }
                    } catch (...) { errs[ti] = std::current_exception(); }
                });
            for (auto& t : th) t.join();
            for (auto& e : errs) if (e) std::rethrow_exception(e);
        }
    }
    for (size_t qi = 0; qi < nq; qi++) {   // serial merge: global term slots and plan numbers
        Prepared& pr = prep[qi];
        const u32 slot_base = nslots;
        const int plan_base = (int)plans.size();
        for (auto& t : pr.terms) t.nfound_slot += (int)slot_base;
        nslots += (u32)pr.terms.size();
        if (pr.status) continue;   // syntax error: its partial plans are dropped
        for (auto& pl : local_plans[qi]) {
            for (auto& in : pl.instr) if (in.op == QOP_COUNT) in.a += slot_base;
            plans.push_back(std::move(pl));
        }
        if (pr.main_plan >= 0) pr.main_plan += plan_base;
        if (pr.field_plan >= 0) pr.field_plan += plan_base;
    }
    ms_plan = since(t_phase);
    // ---- device execution ---------------------------------------------------------------------------
    // Plans run in waves whose scratch (worst-case list areas) fits a memory budget; one kernel launch per wave,
    // one CTA per program.
    const size_t np = plans.size();
    u64 q_launches_pre = 0;
    std::vector<u32> h_nfound(nslots + 1, 0), h_nhits(np + 1, 0), h_npages(np + 1, 0);   // indexed by plan
    std::vector<u32> w_nhits(np + 1, 0), w_npages(np + 1, 0);                            // indexed by launch position
    struct WaveHost {                              // results of one wave, page-locked: the hit download runs at PCIe speed
        PinnedVec<u32> hit_page, page_first;
        PinnedVec<int> hit_pos;
        PinnedVec<float> page_rank;
    };
    std::vector<std::unique_ptr<WaveHost>> wave_host;   // stable addresses: earlier waves are shaped while later ones run
    wave_host.reserve(4 * np + 4);                      // never reallocated while the shaping threads index it (retries included)
    std::vector<u32> plan_wave(np + 1, 0);
    std::vector<u64> hoff(np + 1, 0), poff(np + 1, 0);                                   // by plan: start in its wave's arrays
    u64 hits_total = 0, pages_total = 0;
    for (auto& pl : plans) {   // work estimate (and the same layout the planner made)
        std::vector<u64> caps;
        for (auto& in : pl.instr)
            if (in.op == QOP_LOAD) caps.push_back(in.cap);
        relayout(pl, caps);
    }
    // ---- decoded-list cache -------------------------------------------------------------------------------------------------
    // A Zipf batch asks for the same long lists over and over; a list of at least DCD_SEARCH_CACHE_WORDS ushorts (default
    // 16384) that a single-code LOAD names is decoded ONCE per resident index by all SMs (launch_decode_jobs) and every later
    // LOAD of it is a pointer (QOP_LOADC): no decode, no copy.  DCD_SEARCH_CACHE_MB bounds the cache (default 16384).
    ensure_code_nvals(c, s);
    {
        u64 min_words = 16384, budget_mb = 16384;
        if (const char* e = getenv("DCD_SEARCH_CACHE_WORDS")) min_words = (u64)std::max<long long>(0, atoll(e));
        if (const char* e = getenv("DCD_SEARCH_CACHE_MB")) budget_mb = (u64)std::max<long long>(0, atoll(e));
        ResidentIndex& ix = c->ix;
        std::vector<u32> missing;
        if (budget_mb) {
            std::unordered_set<u32> seen;
            for (auto& pl : plans)
                for (auto& in : pl.instr)
                    if (in.op == QOP_LOAD && in.b == 1) {
                        const u32 code = pl.codes[in.a];
                        if (ix.h_code_off[code + 1] - ix.h_code_off[code] < min_words || ix.dcache.count(code) || !seen.insert(code).second) continue;
                        const u64 bytes = (u64)ix.h_code_nvals[code] * 8;
                        if (ix.dcache_bytes + bytes > (budget_mb << 20)) continue;
                        ix.dcache_bytes += bytes;
                        missing.push_back(code);
                    }
        }
        if (!missing.empty()) {
            const u32 nj = (u32)missing.size();
            u64 total_vals = 0;
            for (u32 code : missing) total_vals += ix.h_code_nvals[code];
            DevBuf<u64> vals(pool, total_vals + 1);
            std::vector<u64> woff(nj), nw(nj), base(nj), outp(nj);
            std::vector<u32> cjob, cidx, jfirst(nj + 1, 0);
            std::vector<u64> hbase;
            if (ix.d_code_base.n) {   // a shard of a sharded build: lists start from the code's first-delta base
                hbase.resize(ix.ncodes);
                DCD_CUDA(cudaMemcpyAsync(hbase.data(), ix.d_code_base.p, (size_t)ix.ncodes * 8, cudaMemcpyDeviceToHost, s));
                DCD_CUDA(cudaStreamSynchronize(s));
            }
            u64 at = 0;
            for (u32 j = 0; j < nj; j++) {
                const u32 code = missing[j];
                woff[j] = ix.h_code_off[code];
                nw[j] = ix.h_code_off[code + 1] - ix.h_code_off[code];
                base[j] = hbase.empty() ? 0 : hbase[code];
                outp[j] = (u64)(uintptr_t)(vals.p + at);
                ix.dcache[code] = {vals.p + at, ix.h_code_nvals[code]};
                at += ix.h_code_nvals[code];
                jfirst[j] = (u32)cjob.size();
                for (u64 k = 0; k * 4096 < nw[j]; k++) { cjob.push_back(j); cidx.push_back((u32)k); }
            }
            jfirst[nj] = (u32)cjob.size();
            const u32 nch = (u32)cjob.size();
            DevBuf<u64> d_woff(pool, nj), d_nw(pool, nj), d_base(pool, nj), d_out(pool, nj), d_psum(pool, nch + 1);
            DevBuf<u32> d_cjob(pool, nch + 1), d_cidx(pool, nch + 1), d_jfirst(pool, nj + 1), d_pcnt(pool, nch + 1);
            DCD_CUDA(cudaMemcpyAsync(d_woff.p, woff.data(), nj * 8, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_nw.p, nw.data(), nj * 8, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_base.p, base.data(), nj * 8, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_out.p, outp.data(), nj * 8, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_cjob.p, cjob.data(), (size_t)nch * 4, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_cidx.p, cidx.data(), (size_t)nch * 4, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_jfirst.p, jfirst.data(), ((size_t)nj + 1) * 4, cudaMemcpyHostToDevice, s));
            DecodeJobs dj{};
            dj.word_off = d_woff.p;
            dj.nwords = d_nw.p;
            dj.base = d_base.p;
            dj.out = reinterpret_cast<u64* const*>(d_out.p);
            dj.chunk_job = d_cjob.p;
            dj.chunk_idx = d_cidx.p;
            dj.job_first_chunk = d_jfirst.p;
            dj.njobs = nj;
            dj.nchunks = nch;
            dj.psum = d_psum.p;
            dj.pcnt = d_pcnt.p;
            launch_decode_jobs(ix.d_arena.p, dj, s);
            DCD_CUDA(cudaStreamSynchronize(s));
            DCD_CUDA(cudaGetLastError());
            ix.dcache_bufs.push_back(std::move(vals));
            q_launches_pre += 3;
            if (trace) fprintf(stderr, "dcd search: decoded-list cache: +%u lists, %.1f M values (cache now %.1f MB)\n", nj, total_vals * 1e-6, ix.dcache_bytes / 1048576.0);
        }
        if (!ix.dcache.empty())
            for (auto& pl : plans) {
                bool touched = false;
                for (auto& in : pl.instr)
                    if (in.op == QOP_LOAD && in.b == 1) {
                        auto it = ix.dcache.find(pl.codes[in.a]);
                        if (it == ix.dcache.end()) continue;
                        in.op = QOP_LOADC;
                        in.src = (u64)(uintptr_t)it->second.first;
                        in.cap = it->second.second;
                        in.lim_cap = in.cap;
                        touched = true;
                    }
                if (touched) {
                    std::vector<u64> caps;
                    for (auto& in : pl.instr)
                        if (in.op == QOP_LOAD || in.op == QOP_LOADC) caps.push_back(in.cap);
                    relayout(pl, caps);
                }
            }
    }
    // longest programs first: one CTA per program, so the tail of a wave is not one long query running alone
    std::vector<u32> order(np);
    for (size_t i = 0; i < np; i++) order[i] = (u32)i;
    std::stable_sort(order.begin(), order.end(), [&](u32 a, u32 b) {
        return plans[a].work > plans[b].work;
    });
    // ---- early termination (QDesc::limit) -------------------------------------------------------------------------------
    // PrepareSearchResult shapes at most MAX_FOUND_PAGES + 1 pages (Search.cs:409), so a program whose lists are long only
    // needs them up to the coordinate where that many pages are found.  Programs start with their lists cut at the end of
    // page 2 x (MAX_FOUND_PAGES + 1); whatever does not reach the page cap inside the trusted prefix runs again at 32 x, then
    // whole.  Eligible: single-code LOADs only, every COUNT directly behind its LOAD (wordInfo.nFound then comes from the
    // per-code value counts), at most 64 terms, lists long enough to matter.  DCD_SEARCH_LIMIT=0 switches it off.
    bool use_limit = !(comm && comm->nranks > 1) && c->ix.npages > 0;
    if (const char* e = getenv("DCD_SEARCH_LIMIT")) use_limit = use_limit && atoi(e) != 0;
    const u64 page_cap = (u64)std::max(0, c->opt.max_found_pages) + 1;
    if (use_limit) engine_fetch_page_ends(c);
    // level 0: the window expected to hold 3 x the page cap (hit pages estimated from the shortest list: an AND is found about
    // as often as its rarest term; phrases and tight proximity far less), level 1: 6 x that window, then everything
    const int LV_FULL = 2;
    auto set_level = [&](Plan& pl, int lv) {
        if (pl.base_instr.empty()) pl.base_instr = pl.instr;
        pl.instr = pl.base_instr;
        u64 min_n = ~0ull;
        for (auto& in : pl.instr)
            if (in.op == QOP_LOAD || in.op == QOP_LOADC) min_n = std::min<u64>(min_n, in.cap);
        const double est = (double)min_n / (pl.rmax < 64 ? 16.0 : 1.0);
        double frac = 3.0 * (double)page_cap / std::max(est, 1.0) * (lv == 0 ? 1.0 : 6.0);
        if (lv >= LV_FULL || frac > 0.5) { lv = LV_FULL; frac = 1.0; }
        pl.level = lv;
        pl.limit = ~0ull;
        if (lv < LV_FULL) {
            const u64 pg = std::min<u64>(c->ix.npages, (u64)(frac * (double)c->ix.npages) + 1);
            pl.limit = c->ix.h_page_end[pg - 1];
        }
        std::vector<u64> caps;
        for (auto& in : pl.instr)
            if (in.op == QOP_LOAD || in.op == QOP_LOADC) {
                u64 cap = in.cap;
                if (lv < LV_FULL) cap = std::min<u64>(cap, (u64)(4.0 * frac * (double)cap) + 16384);
                if (in.op == QOP_LOADC) in.lim_cap = (u32)cap;   // in.cap stays the cached list's length
                caps.push_back(cap);
            }
        relayout(pl, caps);
    };
    if (use_limit) {
        ensure_code_nvals(c, s);
        for (size_t pi = 0; pi < np; pi++) {
            Plan& pl = plans[pi];
            bool ok = pl.has_mul;
            u64 total = 0;
            u32 nl = 0;
            for (size_t i = 0; i < pl.instr.size() && ok; i++) {
                const QInstr& in = pl.instr[i];
                if (in.op == QOP_LOAD || in.op == QOP_LOADC) { ok = in.b <= 1; total += in.cap; nl++; }
                else if (in.op == QOP_COUNT) ok = i > 0 && (pl.instr[i - 1].op == QOP_LOAD || pl.instr[i - 1].op == QOP_LOADC);
            }
            ok = ok && nl <= 64 && total >= (1u << 17);
            if (ok) set_level(pl, 0);
            if (ok && pl.level == LV_FULL) { pl.instr = pl.base_instr; pl.base_instr.clear(); }   // no window worth cutting: a plain program
        }
        std::stable_sort(order.begin(), order.end(), [&](u32 a, u32 b) {
            return plans[a].work > plans[b].work;
        });
    }
    // ---- exact serving from document-range shards: halo pass + exchange ------------------------------------------
    // operator* ignores document boundaries (IndexSequence.cs:218-284): a gap-run can straddle a shard cut.  Every run of every
    // operator is a sub-chain of U = union of the program's leaf lists with gaps <= rmax, so the shards only disagree with the
    // single-index answer inside the chain of U that bridges a cut.  Each rank publishes the tail of its leaf lists from the
    // start of the chain that ends at its last element (when that chain reaches the next shard); the next non-empty rank puts
    // it in front of its own lists (the halo) and answers from the halo's first element on, the publishing rank stops there.
    // Hits a rank finds left of its own coordinate range travel back to the rank holding those pages before hits -> pages.
    const bool exact = comm && comm->nranks > 1;
    struct Exact {
        int R = 1, me = 0;
        std::vector<u32> load_base;            // by plan
        u32 nslots = 0;
        DevBuf<u64> halo_vals, d_halo_off, d_next;
        DevBuf<u32> d_halo_cnt;
        std::vector<u64> keep_below;           // by plan
        u64 range_end = ~0ull;
        // deferred final lists (phase B): per wave dense buffers, per plan pointers
        std::vector<DevBuf<u64>> wave_own, wave_foreign;
        std::vector<u64> own_ptr, foreign_ptr;   // by plan: device addresses
        std::vector<u32> own_n, foreign_n;
        bool phase_b = false;
    } ex;
    if (exact) {
        if (!c->ix.global_codes) throw StatusError(DCD_E_STATE, "sharded search needs the resident index of a sharded build");
        ex.R = comm->nranks;
        ex.me = comm->rank;
        const int R = ex.R, me = ex.me;
        auto tx = now();
        auto lap = [&](const char* what) {
            if (trace) { DCD_CUDA(cudaStreamSynchronize(s)); fprintf(stderr, "dcd search (exact): %s %.1f ms\n", what, since(tx)); tx = now(); }
        };
        ensure_next_first(c, comm, s);
        ex.range_end = me + 1 < R ? c->ix.shard_origins[me + 1] : ~0ull;
        ex.load_base.assign(np + 1, 0);
        for (size_t pi = 0; pi < np; pi++) {
            u32 nl = 0;
            for (auto& in : plans[pi].instr) nl += (in.op == QOP_LOAD || in.op == QOP_LOADC) ? 1u : 0u;
            if (nl > 64) throw StatusError(DCD_E_UNSUPPORTED, "sharded search: more than 64 terms in one request");
            ex.load_base[pi + 1] = ex.load_base[pi] + nl;
        }
        ex.nslots = ex.load_base[np];
        // every rank must have planned the same programs (same batch, same unified dictionary)
        {
            DevBuf<u64> d_meta(pool, 2), d_all(pool, (size_t)2 * R);
            const u64 meta[2] = {(u64)np, (u64)ex.nslots};
            std::vector<u64> all((size_t)2 * R);
            DCD_CUDA(cudaMemcpyAsync(d_meta.p, meta, 16, cudaMemcpyHostToDevice, s));
            comm->allgather(d_meta.p, d_all.p, 16, s);
            DCD_CUDA(cudaMemcpyAsync(all.data(), d_all.p, all.size() * 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            for (int r = 0; r < R; r++)
                if (all[2 * r] != np || all[2 * r + 1] != ex.nslots)
                    throw StatusError(DCD_E_STATE, "sharded search: the ranks were given different batches");
        }
        // ---- phase A: halo pass (LOADs only), waves cut by the scratch budget -------------------------------------------------
        u64 halo_cap = 32ull << 20;   // elements (256 MB); DCD_HALO_MB overrides
        if (const char* e = getenv("DCD_HALO_MB")) halo_cap = std::max<u64>(1, (u64)atoll(e)) << 17;
        DevBuf<u64> my_vals(pool, halo_cap), my_off(pool, (size_t)ex.nslots + 1), d_cursor(pool, 2);
        DevBuf<u32> my_cnt(pool, (size_t)ex.nslots + 1);
        DevBuf<HaloDesc> my_desc(pool, np + 1);
        DCD_CUDA(cudaMemsetAsync(d_cursor.p, 0, 16, s));
        DCD_CUDA(cudaMemsetAsync(my_cnt.p, 0, my_cnt.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(my_off.p, 0, my_off.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(my_desc.p, 0, my_desc.bytes(), s));
        size_t fb = 0, tb = 0;
        cudaMemGetInfo(&fb, &tb);
        const u64 a_budget = std::max<u64>((u64)((fb + pool.bytes_cached) / 4), 128ull << 20);
        auto load_units = [](const Plan& pl, u64* tmp) {
            u64 u = 0;
            *tmp = 1;
            for (auto& in : pl.instr)
                if (in.op == QOP_LOAD) { u += (u64)(in.b > 1 ? 2 : 1) * in.cap; if (in.b > 1) *tmp = std::max<u64>(*tmp, in.cap); }
            return u;   // cached LOADs (QOP_LOADC) are read in place
        };
        for (size_t k0 = 0; k0 < np;) {
            size_t k1 = k0;
            u64 acc = 0;
            while (k1 < np) {
                u64 tmp;
                const u64 b = (load_units(plans[k1], &tmp) + tmp + 2) * 8;
                if (k1 > k0 && (acc + b > a_budget || k1 - k0 >= 4096)) break;
                acc += b;
                k1++;
            }
            const size_t nw = k1 - k0;
            std::vector<QDesc> desc(nw);
            std::vector<QInstr> instr;
            std::vector<u32> codes, lbase(nw);
            std::vector<u64> rmax(nw);
            u64 list_total = 0, tmp_total = 0;
            for (size_t k = 0; k < nw; k++) {
                const Plan& pl = plans[k0 + k];
                QDesc& d = desc[k];
                d.instr_begin = (u32)instr.size();
                const u32 cb = (u32)codes.size();
                for (auto in : pl.instr) {
                    if (in.op != QOP_LOAD && in.op != QOP_LOADC) continue;
                    in.a += cb;
                    instr.push_back(in);
                }
                d.instr_end = (u32)instr.size();
                codes.insert(codes.end(), pl.codes.begin(), pl.codes.end());
                u64 tmp;
                const u64 lu = load_units(pl, &tmp);
                d.list_off = list_total;
                d.tmp_off = tmp_total;
                d.tmp_cap = tmp;
                d.hit_off = 0;
                d.hit_cap = 0;
                list_total += lu + 1;
                tmp_total += tmp + 1;
                lbase[k] = ex.load_base[k0 + k];
                rmax[k] = pl.has_mul ? pl.rmax : 0;
            }
            DevBuf<QDesc> d_desc(pool, nw);
            DevBuf<QInstr> d_instr(pool, instr.size());
            DevBuf<u32> d_codes(pool, codes.size()), d_lbase(pool, nw), d_err(pool, 16);
            DevBuf<u64> d_rmax(pool, nw), d_lists(pool, list_total), d_tmp(pool, tmp_total);
            DCD_CUDA(cudaMemcpyAsync(d_desc.p, desc.data(), nw * sizeof(QDesc), cudaMemcpyHostToDevice, s));
            if (!instr.empty()) DCD_CUDA(cudaMemcpyAsync(d_instr.p, instr.data(), instr.size() * sizeof(QInstr), cudaMemcpyHostToDevice, s));
            if (!codes.empty()) DCD_CUDA(cudaMemcpyAsync(d_codes.p, codes.data(), codes.size() * 4, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_lbase.p, lbase.data(), nw * 4, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemcpyAsync(d_rmax.p, rmax.data(), nw * 8, cudaMemcpyHostToDevice, s));
            DCD_CUDA(cudaMemsetAsync(d_err.p, 0, 64, s));
            QBatch qb{};
            qb.desc = d_desc.p;
            qb.instr = d_instr.p;
            qb.code_ids = d_codes.p;
            qb.nqueries = (u32)nw;
            qb.arena = c->ix.d_arena.p;
            qb.code_off = c->ix.d_code_off.p;
            qb.code_base = c->ix.d_code_base.n ? c->ix.d_code_base.p : nullptr;
            qb.coord_origin = c->ix.coord_origin;
            qb.lists = d_lists.p;
            qb.tmp_dec = d_tmp.p;
            qb.error = d_err.p;
            qb.load_base = d_lbase.p;
            HaloPass hp{};
            hp.rmax = d_rmax.p;
            hp.next_first = c->ix.d_next_first.p;
            hp.desc = my_desc.p + k0;
            hp.slot_cnt = my_cnt.p;
            hp.slot_off = my_off.p;
            hp.vals = my_vals.p;
            hp.vals_cap = halo_cap;
            hp.cursor = reinterpret_cast<unsigned long long*>(d_cursor.p);
            launch_query_halo(qb, hp, s);
            u32 herr[16];
            DCD_CUDA(cudaMemcpyAsync(herr, d_err.p, 64, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            DCD_CUDA(cudaGetLastError());
            if (herr[0] == 3) throw StatusError(DCD_E_CAPACITY, "sharded search: the halo buffer is full (raise DCD_HALO_MB)");
            if (herr[0]) throw StatusError(DCD_E_CAPACITY, "sharded search: halo pass exceeded its planned capacity");
            q_launches_pre++;
            k0 = k1;
        }
        lap("halo pass");
        // ---- exchange: descriptors of every rank, slot counts / offsets, payloads ------------------------------------------------
        u64 my_total = 0;
        DCD_CUDA(cudaMemcpyAsync(&my_total, d_cursor.p, 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DevBuf<u64> d_tot(pool, 1), d_tot_all(pool, R);
        std::vector<u64> tot_all(R);
        DCD_CUDA(cudaMemcpyAsync(d_tot.p, &my_total, 8, cudaMemcpyHostToDevice, s));
        comm->allgather(d_tot.p, d_tot_all.p, 8, s);
        DCD_CUDA(cudaMemcpyAsync(tot_all.data(), d_tot_all.p, (size_t)R * 8, cudaMemcpyDeviceToHost, s));
        DevBuf<HaloDesc> all_desc(pool, (np + 1) * R);
        DevBuf<u32> all_cnt(pool, ((size_t)ex.nslots + 1) * R);
        DevBuf<u64> all_off(pool, ((size_t)ex.nslots + 1) * R);
        comm->allgather(my_desc.p, all_desc.p, (np + 1) * sizeof(HaloDesc), s);
        comm->allgather(my_cnt.p, all_cnt.p, ((size_t)ex.nslots + 1) * 4, s);
        comm->allgather(my_off.p, all_off.p, ((size_t)ex.nslots + 1) * 8, s);
        DCD_CUDA(cudaStreamSynchronize(s));
        std::vector<size_t> vcnt(R), voff(R);
        size_t vtot = 0;
        for (int r = 0; r < R; r++) { vcnt[r] = (size_t)tot_all[r] * 8; voff[r] = vtot; vtot += (vcnt[r] + 15) & ~size_t(15); }
        DevBuf<u8> all_vals(pool, vtot + 16);
        comm->allgatherv(my_vals.p, all_vals.p, vcnt.data(), voff.data(), s);
        std::vector<HaloDesc> hd((np + 1) * R);
        std::vector<u32> hcnt(((size_t)ex.nslots + 1) * R);
        std::vector<u64> hoff2(((size_t)ex.nslots + 1) * R);
        DCD_CUDA(cudaMemcpyAsync(hd.data(), all_desc.p, hd.size() * sizeof(HaloDesc), cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaMemcpyAsync(hcnt.data(), all_cnt.p, hcnt.size() * 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaMemcpyAsync(hoff2.data(), all_off.p, hoff2.size() * 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        my_vals.reset();
        lap("halo exchange");
        // ---- compose the left halo of every program: walk the ranks to the left while their chains reach this one --------------------
        auto D = [&](int r, size_t pi) -> const HaloDesc& { return hd[(size_t)r * (np + 1) + pi]; };
        std::vector<u32> h_cnt((size_t)ex.nslots + 1, 0);
        std::vector<u64> h_off((size_t)ex.nslots + 1, 0);
        std::vector<u64> ssrc, sdst;
        std::vector<u32> scnt;
        ex.keep_below.assign(np + 1, ~0ull);
        u64 composed = 0;
        for (size_t pi = 0; pi < np; pi++) {
            const HaloDesc& mine = D(me, pi);
            if (!(mine.flags & 1u)) continue;   // nothing of U here: this rank contributes no hits for the program
            std::vector<int> contrib;
            for (int r = me - 1; r >= 0; r--) {
                const HaloDesc& h = D(r, pi);
                if (!(h.flags & 1u)) continue;
                if (!(h.flags & 2u)) break;
                contrib.insert(contrib.begin(), r);
                if (!(h.flags & 4u)) break;
            }
            const u32 s0 = ex.load_base[pi], s1 = ex.load_base[pi + 1];
            for (u32 sl = s0; sl < s1; sl++) {
                h_off[sl] = composed;
                for (int r : contrib) {
                    const u32 cnt = hcnt[(size_t)r * (ex.nslots + 1) + sl];
                    if (!cnt) continue;
                    ssrc.push_back(voff[r] / 8 + hoff2[(size_t)r * (ex.nslots + 1) + sl]);
                    sdst.push_back(composed);
                    scnt.push_back(cnt);
                    h_cnt[sl] += cnt;
                    composed += cnt;
                }
            }
            if (mine.flags & 2u) ex.keep_below[pi] = ((mine.flags & 4u) && !contrib.empty()) ? D(contrib.front(), pi).lo : mine.lo;
        }
        ex.halo_vals = DevBuf<u64>(pool, composed + 1);
        copy_segments(pool, reinterpret_cast<const u64*>(all_vals.p), ex.halo_vals.p, ssrc, sdst, scnt, s);
        ex.d_halo_cnt = DevBuf<u32>(pool, (size_t)ex.nslots + 1);
        ex.d_halo_off = DevBuf<u64>(pool, (size_t)ex.nslots + 1);
        DCD_CUDA(cudaMemcpyAsync(ex.d_halo_cnt.p, h_cnt.data(), h_cnt.size() * 4, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(ex.d_halo_off.p, h_off.data(), h_off.size() * 8, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        // ---- re-lay every program's list area with room for the halos ----------------------------------------------------------
        for (size_t pi = 0; pi < np; pi++) {
            Plan& pl = plans[pi];
            std::vector<u64> caps;
            std::vector<u8> copy;
            u32 sl = ex.load_base[pi];
            for (auto& in : pl.instr)
                if (in.op == QOP_LOAD || in.op == QOP_LOADC) { copy.push_back(h_cnt[sl] ? 1 : 0); caps.push_back((u64)in.cap + h_cnt[sl++]); }
            relayout(pl, caps, &copy);
        }
        lap("compose + re-layout");
        ex.own_ptr.assign(np + 1, 0);
        ex.foreign_ptr.assign(np + 1, 0);
        ex.own_n.assign(np + 1, 0);
        ex.foreign_n.assign(np + 1, 0);
        ex.phase_b = true;
    }
    // ---- result shaping (PrepareSearchResult, Search.cs:365-420) -----------------------------------------
    struct Shaped {
        std::vector<HitPage> pages;
        std::vector<int32_t> pos;
        struct D { i64 doc; u16s name; int first; int n; bool found; };
        std::vector<D> docs;
    };
    // Index made of bulk segments only: document names are "<source>:d<number>", unique per ordinal, and no page has the
    // id "0" -- documents are told apart by ordinal and no per-page strings are built (names only when a filter needs them).
    bool all_bulk = true;
    for (auto& sg : c->ix.segments) {
        if (!sg.bulk) { all_bulk = false; break; }
        engine_fetch_segment_meta(sg);
    }
    engine_fetch_page_ends(c);
    // -filter: patterns (Search.cs:389-395: Regex.Match on the document name).  Compiled once per request: ECMAScript syntax
    // over the UTF-8 name (the common subset of .NET's syntax: literals, classes, quantifiers, groups, alternation); a
    // pattern std::regex cannot compile is matched as a plain substring (INTEGRATION.md lists the deviation).
    struct NameFilter { std::string pat; bool compiled = false; std::regex re; };
    auto compile_filters = [](const std::vector<u16s>& filter) {
        std::vector<NameFilter> out(filter.size());
        for (size_t i = 0; i < filter.size(); i++) {
            out[i].pat = utf16_to_utf8(filter[i]);
            try { out[i].re = std::regex(out[i].pat); out[i].compiled = true; } catch (const std::regex_error&) {}
        }
        return out;
    };
    auto name_matches = [](const std::vector<NameFilter>& fs, const u16s& name16) {
        const std::string name = utf16_to_utf8(name16);
        for (const auto& f : fs)
            if (f.compiled ? std::regex_search(name, f.re) : name.find(f.pat) != std::string::npos) return true;
        return false;
    };
    auto shape_bulk = [&](int plan, const std::vector<u16s>& filter) {
        Shaped sh;
        const std::vector<NameFilter> nf = compile_filters(filter);
        const u32 nh = h_nhits[plan], npg = h_npages[plan];
        const u64 ho = hoff[plan], po = poff[plan];
        const WaveHost& W = *wave_host[plan_wave[plan]];
        const int maxp = c->opt.max_found_pages, maxd = c->opt.max_found_docs;
        int nfound_docs = 0;
        i64 prev_doc = -1;
        size_t si = 0;          // pages ascend: segment and document cursors only move forward
        u64 dcur = 0;
        auto& segs = c->ix.segments;
        sh.pages.reserve(std::min<u32>(npg, (u32)maxp + 1));
        sh.pos.reserve(nh);
        for (u32 k = 0; k < npg; k++) {
            const u32 a = W.page_first[po + k], e = k + 1 < npg ? W.page_first[po + k + 1] : nh;
            if (W.hit_page[ho + a] == 0xFFFFFFFFu) break;
            HitPage hp;
            hp.ordinal = W.hit_page[ho + a];
            hp.rank = W.page_rank[po + k];
            hp.first = (u32)sh.pos.size();
            hp.npos = e - a;
            const bool last = (int)sh.pages.size() + 1 > maxp;
            if (last) { hp.npos = 1; hp.rank = 1.0f; }
            sh.pos.insert(sh.pos.end(), W.hit_pos.data() + ho + a, W.hit_pos.data() + ho + a + hp.npos);
            if (nfound_docs >= maxd && prev_doc == -2) {   // MAX_FOUND_DOCS reached and the last found document is behind: no later
                sh.pages.push_back(hp);                    // page can add a document, only the page list grows (Search.cs:385-409)
                if (last) break;
                continue;
            }
            const u64 ord = (u64)hp.ordinal;
            while (si < segs.size() && ord >= segs[si].first_ordinal + segs[si].count) { si++; dcur = 0; }
            if (si >= segs.size()) break;
            PageSegment& sg = segs[si];
            const u64 kk = ord - sg.first_ordinal;
            const u64 orig = sg.identity ? kk : sg.h_page_orig[kk];
            for (int step = 0; step < 4 && dcur + 1 < sg.ndocs && sg.h_doc_first_page[dcur + 1] <= orig; step++) dcur++;
            if (dcur + 1 < sg.ndocs && sg.h_doc_first_page[dcur + 1] <= orig) {   // far jump: last d with first_page[d] <= orig
                auto fp = sg.h_doc_first_page.begin();
                dcur = (u64)(std::upper_bound(fp + dcur + 1, fp + sg.ndocs, orig) - fp) - 1;
            }
            const i64 doc = (i64)(sg.first_doc + dcur);
            if (doc != prev_doc) {
                Shaped::D d;
                d.doc = doc;
                d.first = (int)sh.pages.size();
                d.n = 0;
                d.found = false;
                if (nfound_docs < maxd) {
                    bool matched = filter.empty();
                    if (!matched) {
                        d.name = engine_doc_name(c, (u64)doc);
                        matched = name_matches(nf, d.name);
                    }
                    if (matched) { d.found = true; nfound_docs++; }
                } else {
                    prev_doc = -2;   // first page of a document past the cap: the document bookkeeping ends here
                    sh.pages.push_back(hp);
                    if (last) break;
                    continue;
                }
                sh.docs.push_back(std::move(d));
                prev_doc = doc;
            }
            sh.docs.back().n++;
            sh.pages.push_back(hp);
            if (last) break;
        }
        return sh;
    };
    auto shape_general = [&](int plan, const std::vector<u16s>& filter) {
        Shaped sh;
        const std::vector<NameFilter> nf = compile_filters(filter);
        const u32 nh = h_nhits[plan], npg = h_npages[plan];
        const u64 ho = hoff[plan], po = poff[plan];
        const WaveHost& W = *wave_host[plan_wave[plan]];
        const int maxp = c->opt.max_found_pages, maxd = c->opt.max_found_docs;
        std::unordered_set<u16s> found_names;
        i64 prev_doc = -1;
        for (u32 k = 0; k < npg; k++) {
            const u32 a = W.page_first[po + k], e = k + 1 < npg ? W.page_first[po + k + 1] : nh;
            if (W.hit_page[ho + a] == 0xFFFFFFFFu) break;   // coordinate beyond the last page (malformed index)
            HitPage hp;
            hp.ordinal = W.hit_page[ho + a];
            hp.rank = W.page_rank[po + k];
            hp.first = (u32)sh.pos.size();
            hp.npos = e - a;
            const bool last = (int)sh.pages.size() + 1 > maxp;   // Search.cs:409: break right after this page's first position
            if (last) { hp.npos = 1; hp.rank = 1.0f; }
            for (u32 i = 0; i < hp.npos; i++) sh.pos.push_back(W.hit_pos[ho + a + i]);
            u16s id;
            u64 doc, st, en;
            engine_page_info(c, (u64)hp.ordinal, &id, &doc, &st, &en);
            if ((i64)doc != prev_doc) {
                Shaped::D d;
                d.doc = (i64)doc;
                d.name = engine_doc_name(c, doc);
                d.first = (int)sh.pages.size();
                d.n = 0;
                d.found = false;
                if ((int)found_names.size() < maxd) {
                    const bool matched = filter.empty() || name_matches(nf, d.name);
                    if (matched && found_names.insert(d.name).second) d.found = true;
                }
                sh.docs.push_back(d);
                prev_doc = (i64)doc;
            }
            sh.docs.back().n++;
            sh.pages.push_back(hp);
            if (last) break;
        }
        return sh;
    };
    auto shape = [&](int plan, const std::vector<u16s>& filter) { return all_bulk ? shape_bulk(plan, filter) : shape_general(plan, filter); };

    auto finish_query = [&](size_t qi) {
        Prepared& pr = prep[qi];
        QueryResult& r = out->q[qi];
        r.status = pr.status;
        if (pr.status) return;
        int plan = pr.main_plan >= 0 ? pr.main_plan : pr.field_plan;   // Search.cs:538
        if (plan < 0) return;                                          // empty SearchResult, Search.cs:615-616
        Shaped sh = shape(plan, pr.filter);
        std::unordered_set<u16s> fnames;
        const bool have_f = pr.field_plan >= 0;
        std::unordered_set<i64> fdocs;                                 // all-bulk index: names are unique per ordinal
        if (have_f) {
            Shaped sf = shape(pr.field_plan, {});
            for (auto& d : sf.docs) if (d.found) { if (all_bulk) fdocs.insert(d.doc); else fnames.insert(d.name); }
        }
        r.pages = std::move(sh.pages);
        r.pos = std::move(sh.pos);
        r.docs.reserve(sh.docs.size());
        for (auto& d : sh.docs) {
            if (!d.found) continue;
            if (have_f && !(all_bulk ? fdocs.count(d.doc) != 0 : fnames.count(d.name) != 0)) continue;   // CombineSearchResults, Search.cs:423-428
            HitDoc hd;
            hd.doc = d.doc;
            hd.first_page = d.first;
            hd.npages = d.n;
            double sum = 0;                                            // Enumerable.Sum(float) accumulates in double
            for (int k = 0; k < d.n; k++) sum += (double)r.pages[d.first + k].rank;
            float rank = (float)sum;
            rank = 1 + (float)std::log((double)rank);                  // Search.cs:555
            if (!all_bulk) {                                           // bulk pages are numbered from "1"
                u16s id;
                u64 doc, st, en;
                engine_page_info(c, (u64)r.pages[d.first].ordinal, &id, &doc, &st, &en);
                if (id == u"0") rank *= 10.0f;                         // DOC_RANK_MULTIPLY, Index.cs:115; Search.cs:556-557
            }
            hd.rank = rank;
            r.docs.push_back(hd);
        }
        std::stable_sort(r.docs.begin(), r.docs.end(), [](const HitDoc& a, const HitDoc& b) { return a.rank < b.rank; });   // Search.cs:599
    };
    // Shaping of a request starts as soon as the waves holding its programs have been downloaded: worker threads shape
    // earlier waves while the device runs the later ones (bulk-only indexes; the general path mutates lazily fetched page
    // metadata and is shaped serially at the end).
    std::vector<int> pending(nq, 0);
    std::vector<u32> plan_req(np + 1, 0);
    for (size_t qi = 0; qi < nq; qi++) {
        if (prep[qi].status) continue;
        if (prep[qi].main_plan >= 0) { pending[qi]++; plan_req[prep[qi].main_plan] = (u32)qi; }
        if (prep[qi].field_plan >= 0) { pending[qi]++; plan_req[prep[qi].field_plan] = (u32)qi; }
    }
    std::mutex rq_mu;
    std::condition_variable rq_cv;
    std::vector<size_t> ready;
    bool rq_closed = false;
    std::vector<std::thread> workers;
    std::vector<std::exception_ptr> worker_err;
    unsigned worker_cap = 32;
    if (const char* e = getenv("DCD_SEARCH_WORKERS")) worker_cap = (unsigned)std::max(1, atoi(e));
    const unsigned nworkers = (all_bulk && nq >= 64 && !sink) ? std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), worker_cap) : 0u;
    if (trace) fprintf(stderr, "dcd search: %u shaping threads (%u hardware threads)\n", nworkers, std::thread::hardware_concurrency());
    worker_err.resize(nworkers);
    for (unsigned wi = 0; wi < nworkers; wi++)
        workers.emplace_back([&, wi] {
            try {
                for (;;) {
                    std::vector<size_t> mine;
                    {
                        std::unique_lock<std::mutex> lk(rq_mu);
                        rq_cv.wait(lk, [&] { return !ready.empty() || rq_closed; });
                        if (ready.empty()) return;
                        for (int k = 0; k < 16 && !ready.empty(); k++) { mine.push_back(ready.back()); ready.pop_back(); }
                    }
                    for (size_t qi : mine) finish_query(qi);
                }
            } catch (...) { worker_err[wi] = std::current_exception(); }
        });
    struct JoinWorkers {   // also on the exception path
        std::mutex& mu; std::condition_variable& cv; bool& closed; std::vector<std::thread>& th;
        void close() { { std::lock_guard<std::mutex> lk(mu); closed = true; } cv.notify_all(); for (auto& t : th) if (t.joinable()) t.join(); }
        ~JoinWorkers() { close(); }
    } join_workers{rq_mu, rq_cv, rq_closed, workers};
    double q_ms = 0;
    u64 q_bytes = 0, q_launches = 0;
    size_t free_b = 0, total_b = 0;
    cudaMemGetInfo(&free_b, &total_b);
    // Waves overlap on the device: up to `depth` waves are in flight, each on its own stream, so the tail of one wave (a
    // few long programs, one CTA each) runs beside the next waves' programs instead of beside idle SMs; the host packs
    // and downloads wave k while waves k+1.. compute.  The memory budget is shared by the waves in flight.
    int depth = 4;
    if (const char* e = getenv("DCD_SEARCH_INFLIGHT")) depth = std::max(1, std::min(8, atoi(e)));
    size_t max_wave_programs = 512;
    if (const char* e = getenv("DCD_SEARCH_WAVE_PROGRAMS")) max_wave_programs = (size_t)std::max(1, atoi(e));
    u64 budget = std::max<u64>((u64)((free_b + pool.bytes_cached) / 2), 256ull << 20);
    if (const char* e = getenv("DCD_SEARCH_WAVE_KB")) budget = std::max<u64>(1, (u64)atoll(e)) << 10;   // tests: many small waves
    else budget /= (u64)depth;
    auto plan_bytes = [](const Plan& pl) { return (pl.list_units + 1) * 8 + (pl.tmp_cap + 1) * 9 + (pl.hit_cap + 1) * 16; };
    DevBuf<u32> d_nfound(pool, nslots + 1);
    if (np) DCD_CUDA(cudaMemsetAsync(d_nfound.p, 0, d_nfound.bytes(), s));
    struct WaveDev {   // one wave in flight: its device buffers live until its results are on the host
        size_t k0 = 0, nw = 0;
        std::vector<QDesc> desc;
        std::vector<QInstr> instr;
        std::vector<u32> codes;
        DevBuf<QDesc> d_desc;
        DevBuf<QInstr> d_instr;
        DevBuf<u32> d_codes, d_hit_page, d_page_first, d_nhits, d_npages, d_err;
        DevBuf<u64> d_lists, d_tmp_dec, d_cta_ns, d_keep, d_fin_ptr;
        DevBuf<u32> d_lbase, d_fin_n, d_fin_foreign, d_incomplete;
        std::vector<u32> h_incomplete;
        std::vector<u32> h_lbase;
        std::vector<u64> h_keep;
        bool pages_pass = false;   // exact sharded serving, phase D: this wave ran k_query_pages
        DevBuf<u8> d_tag;
        DevBuf<int> d_hit_pos;
        DevBuf<float> d_page_rank;
        QBatch qb{};
        cudaStream_t st = nullptr;
        cudaEvent_t ev1 = nullptr;
        ~WaveDev() { if (ev1) cudaEventDestroy(ev1); }
    };
    std::deque<std::unique_ptr<WaveDev>> inflight;
    struct WaveStreams {   // declared after `inflight`: drained before the waves' buffers go back to the pool, also on the exception path
        std::vector<cudaStream_t> st;
        cudaEvent_t begin = nullptr, first = nullptr;
        ~WaveStreams() {
            for (auto x : st) { cudaStreamSynchronize(x); cudaStreamDestroy(x); }
            if (begin) cudaEventDestroy(begin);
            if (first) cudaEventDestroy(first);
        }
    } ws;
    if (np) {
        DCD_CUDA(cudaEventCreateWithFlags(&ws.begin, cudaEventDisableTiming));
        DCD_CUDA(cudaEventCreate(&ws.first));
        DCD_CUDA(cudaEventRecord(ws.begin, s));   // the index, d_nfound and recycled pool memory were last touched on s
        for (int i = 0; i < depth; i++) {
            cudaStream_t x;
            DCD_CUDA(cudaStreamCreateWithFlags(&x, cudaStreamNonBlocking));
            ws.st.push_back(x);
            DCD_CUDA(cudaStreamWaitEvent(x, ws.begin, 0));
        }
    }
    std::vector<u64> plan_pairs(sink ? np + 1 : 0, 0);   // pages-only mode: device address of every plan's pairs
    std::vector<u32> retry;                              // early termination: programs to run again one level up
    u32 gallop_ratio = 12;
    if (const char* e = getenv("DCD_SEARCH_GALLOP")) gallop_ratio = (u32)std::max(0, atoi(e));
    size_t waves_launched = 0;
    auto launch_wave = [&](size_t k0, size_t k1) {
        t_phase = now();
        std::unique_ptr<WaveDev> wp(new WaveDev);
        WaveDev& W = *wp;
        const size_t nw = k1 - k0;
        W.k0 = k0;
        W.nw = nw;
        W.st = ws.st[waves_launched % ws.st.size()];
        W.desc.resize(nw);
        u64 list_total = 0, tmp_total = 0, hit_total = 0;
        for (size_t k = 0; k < nw; k++) {
            Plan& pl = plans[order[k0 + k]];
            QDesc& d = W.desc[k];
            d.instr_begin = (u32)W.instr.size();
            const u32 code_base = (u32)W.codes.size();
            for (auto in : pl.instr) {
                if (in.op == QOP_LOAD || in.op == QOP_LOADC) in.a += code_base;
                W.instr.push_back(in);
            }
            d.instr_end = (u32)W.instr.size();
            W.codes.insert(W.codes.end(), pl.codes.begin(), pl.codes.end());
            d.list_off = list_total;
            d.tmp_off = tmp_total;
            d.tmp_cap = pl.tmp_cap;
            d.hit_off = hit_total;
            d.hit_cap = pl.hit_cap;
            d.limit = pl.limit;
            d.rmax = pl.rmax;
            list_total += pl.list_units + 1;
            tmp_total += pl.tmp_cap + 1;
            hit_total += pl.hit_cap + 1;
            if (pl.tmp_cap > 0x7FFFFFFFull) throw StatusError(DCD_E_CAPACITY, "posting lists of one query exceed 2^31 entries");
        }
        W.d_desc = DevBuf<QDesc>(pool, nw);
        W.d_instr = DevBuf<QInstr>(pool, W.instr.size());
        W.d_codes = DevBuf<u32>(pool, W.codes.size());
        W.d_lists = DevBuf<u64>(pool, list_total);
        W.d_tmp_dec = DevBuf<u64>(pool, tmp_total);
        W.d_tag = DevBuf<u8>(pool, tmp_total);
        W.d_hit_page = DevBuf<u32>(pool, hit_total);
        W.d_page_first = DevBuf<u32>(pool, hit_total);
        W.d_nhits = DevBuf<u32>(pool, nw);
        W.d_npages = DevBuf<u32>(pool, nw);
        W.d_err = DevBuf<u32>(pool, 16);
        W.d_hit_pos = DevBuf<int>(pool, hit_total);
        W.d_page_rank = DevBuf<float>(pool, hit_total);
        W.d_cta_ns = DevBuf<u64>(pool, trace ? nw : 0);
        W.d_incomplete = DevBuf<u32>(pool, nw);
        cudaStream_t st = W.st;
        DCD_CUDA(cudaMemsetAsync(W.d_incomplete.p, 0, nw * 4, st));
        DCD_CUDA(cudaMemcpyAsync(W.d_desc.p, W.desc.data(), nw * sizeof(QDesc), cudaMemcpyHostToDevice, st));
        DCD_CUDA(cudaMemcpyAsync(W.d_instr.p, W.instr.data(), W.instr.size() * sizeof(QInstr), cudaMemcpyHostToDevice, st));
        if (!W.codes.empty()) DCD_CUDA(cudaMemcpyAsync(W.d_codes.p, W.codes.data(), W.codes.size() * 4, cudaMemcpyHostToDevice, st));
        DCD_CUDA(cudaMemsetAsync(W.d_err.p, 0, 64, st));
        QBatch& qb = W.qb;
        qb.desc = W.d_desc.p;
        qb.instr = W.d_instr.p;
        qb.code_ids = W.d_codes.p;
        qb.nqueries = (u32)nw;
        qb.arena = c->ix.d_arena.p;
        qb.code_off = c->ix.d_code_off.p;
        qb.code_base = c->ix.d_code_base.n ? c->ix.d_code_base.p : nullptr;
        qb.page_end = c->ix.d_page_end.p;
        qb.npages = c->ix.npages;
        qb.coord_origin = c->ix.coord_origin;
        qb.max_pages = (u32)std::max(0, c->opt.max_found_pages);
        qb.lists = W.d_lists.p;
        qb.tmp_dec = W.d_tmp_dec.p;
        qb.tmp_tag = W.d_tag.p;
        qb.nfound = d_nfound.p;
        qb.hit_page = W.d_hit_page.p;
        qb.hit_pos = W.d_hit_pos.p;
        qb.page_first = W.d_page_first.p;
        qb.page_rank = W.d_page_rank.p;
        qb.nhits = W.d_nhits.p;
        qb.npages_found = W.d_npages.p;
        qb.error = W.d_err.p;
        qb.incomplete = W.d_incomplete.p;
        qb.gallop_ratio = gallop_ratio;
        qb.cta_ns = trace ? W.d_cta_ns.p : nullptr;
        if (ex.phase_b) {   // exact sharded serving: halos in front of the LOAD results, final lists deferred to the exchange
            std::vector<u32>& lbase = W.h_lbase;
            std::vector<u64>& keep = W.h_keep;
            lbase.resize(nw);
            keep.resize(nw);
            for (size_t k = 0; k < nw; k++) { lbase[k] = ex.load_base[order[k0 + k]]; keep[k] = ex.keep_below[order[k0 + k]]; }
            W.d_lbase = DevBuf<u32>(pool, nw);
            W.d_keep = DevBuf<u64>(pool, nw);
            W.d_fin_ptr = DevBuf<u64>(pool, nw);
            W.d_fin_n = DevBuf<u32>(pool, nw);
            W.d_fin_foreign = DevBuf<u32>(pool, nw);
            DCD_CUDA(cudaMemcpyAsync(W.d_lbase.p, lbase.data(), nw * 4, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemcpyAsync(W.d_keep.p, keep.data(), nw * 8, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemsetAsync(W.d_fin_n.p, 0, nw * 4, st));
            DCD_CUDA(cudaMemsetAsync(W.d_fin_foreign.p, 0, nw * 4, st));
            DCD_CUDA(cudaMemsetAsync(W.d_fin_ptr.p, 0, nw * 8, st));
            qb.load_base = W.d_lbase.p;
            qb.halo_cnt = ex.d_halo_cnt.p;
            qb.halo_off = ex.d_halo_off.p;
            qb.halo_vals = ex.halo_vals.p;
            qb.keep_below = W.d_keep.p;
            qb.defer_pages = 1;
            qb.fin_ptr = reinterpret_cast<u64**>(W.d_fin_ptr.p);
            qb.fin_n = W.d_fin_n.p;
            qb.fin_foreign = W.d_fin_foreign.p;
        }
        if (waves_launched == 0) DCD_CUDA(cudaEventRecord(ws.first, st));
        launch_query_batch(qb, st);
        DCD_CUDA(cudaEventCreate(&W.ev1));
        DCD_CUDA(cudaEventRecord(W.ev1, st));
        waves_launched++;
        ms_wave_prep += since(t_phase);
        return wp;
    };
    auto finish_wave = [&](WaveDev& W) {
        t_phase = now();
        const size_t k0 = W.k0, nw = W.nw;
        cudaStream_t st = W.st;
        u32 herr[16];
        DCD_CUDA(cudaMemcpyAsync(w_nhits.data() + k0, W.d_nhits.p, nw * 4, cudaMemcpyDeviceToHost, st));
        DCD_CUDA(cudaMemcpyAsync(w_npages.data() + k0, W.d_npages.p, nw * 4, cudaMemcpyDeviceToHost, st));
        DCD_CUDA(cudaMemcpyAsync(herr, W.d_err.p, 64, cudaMemcpyDeviceToHost, st));
        W.h_incomplete.assign(nw, 0);
        if (W.d_incomplete.n) DCD_CUDA(cudaMemcpyAsync(W.h_incomplete.data(), W.d_incomplete.p, nw * 4, cudaMemcpyDeviceToHost, st));
        DCD_CUDA(cudaStreamSynchronize(st));
        for (size_t k = 0; k < nw; k++)
            if (W.h_incomplete[k]) {   // not enough trusted pages inside the limit: nothing of this run is kept, the program runs again
                w_nhits[k0 + k] = 0;
                w_npages[k0 + k] = 0;
                retry.push_back(order[k0 + k]);
            }
        DCD_CUDA(cudaGetLastError());
        if (herr[0]) throw StatusError(DCD_E_CAPACITY, "query program exceeded its planned capacity");
        ms_kernel_wait += since(t_phase);
        if (ex.phase_b && !W.pages_pass) {   // keep the deferred final lists: own part and the part left of this shard's range
            std::vector<u32> fn(nw), ff(nw);
            DCD_CUDA(cudaMemcpyAsync(fn.data(), W.d_fin_n.p, nw * 4, cudaMemcpyDeviceToHost, st));
            DCD_CUDA(cudaMemcpyAsync(ff.data(), W.d_fin_foreign.p, nw * 4, cudaMemcpyDeviceToHost, st));
            DCD_CUDA(cudaStreamSynchronize(st));
            std::vector<u64> ooff(nw + 1, 0), foff(nw + 1, 0);
            for (size_t k = 0; k < nw; k++) { ooff[k + 1] = ooff[k] + (fn[k] - ff[k]); foff[k + 1] = foff[k] + ff[k]; }
            DevBuf<u64> own(pool, ooff[nw] + 1), foreign(pool, foff[nw] + 1), d_ooff(pool, nw + 1), d_foff(pool, nw + 1);
            DCD_CUDA(cudaMemcpyAsync(d_ooff.p, ooff.data(), (nw + 1) * 8, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemcpyAsync(d_foff.p, foff.data(), (nw + 1) * 8, cudaMemcpyHostToDevice, st));
            launch_pack_final(W.qb, d_ooff.p, d_foff.p, own.p, foreign.p, st);
            DCD_CUDA(cudaStreamSynchronize(st));
            DCD_CUDA(cudaGetLastError());
            for (size_t k = 0; k < nw; k++) {
                const u32 pl = order[k0 + k];
                ex.own_ptr[pl] = (u64)(uintptr_t)(own.p + ooff[k]);
                ex.own_n[pl] = fn[k] - ff[k];
                ex.foreign_ptr[pl] = (u64)(uintptr_t)(foreign.p + foff[k]);
                ex.foreign_n[pl] = ff[k];
            }
            ex.wave_own.push_back(std::move(own));
            ex.wave_foreign.push_back(std::move(foreign));
            float end_ms = 0;
            cudaEventElapsedTime(&end_ms, ws.first, W.ev1);
            q_ms = std::max(q_ms, (double)end_ms);
            q_launches += 2;
            for (u32 id : W.codes) q_bytes += 2 * (c->ix.h_code_off[id + 1] - c->ix.h_code_off[id]);
            return;
        }
        float end_ms = 0;
        cudaEventElapsedTime(&end_ms, ws.first, W.ev1);   // device time of the batch = first wave's start .. last wave's end
        q_ms = std::max(q_ms, (double)end_ms);
        if (trace) {
            std::vector<u64> ns(nw);
            DCD_CUDA(cudaMemcpy(ns.data(), W.d_cta_ns.p, nw * 8, cudaMemcpyDeviceToHost));
            u64 mx = 0, sum = 0;
            for (u64 v : ns) { mx = std::max(mx, v); sum += v; }
            fprintf(stderr, "dcd search: wave of %zu programs ended %.1f ms after the first wave began: longest program %.2f ms, sum of "
                            "program times %.1f ms, first (largest planned) %.2f ms\n", nw, end_ms, mx * 1e-6, sum * 1e-6, ns[0] * 1e-6);
            if (trace > 1) {   // the wave's longest programs with their plan: input of the launch-order cost model
                std::vector<u32> idx(nw);
                for (size_t k = 0; k < nw; k++) idx[k] = (u32)k;
                const size_t top = std::min<size_t>(nw, 6);
                std::partial_sort(idx.begin(), idx.begin() + top, idx.end(), [&](u32 a, u32 b) { return ns[a] > ns[b]; });
                for (size_t t = 0; t < top; t++) {
                    const Plan& pl = plans[order[k0 + idx[t]]];
                    std::string prog;
                    for (const auto& in : pl.instr) {
                        char buf[48];
                        if (in.op == QOP_LOAD) snprintf(buf, sizeof buf, "L%u:%u ", in.b, in.cap);
                        else if (in.op == QOP_MUL) snprintf(buf, sizeof buf, "M%u%s ", in.a, in.b ? "o" : "");
                        else if (in.op == QOP_ADD) snprintf(buf, sizeof buf, "A%u ", in.a);
                        else buf[0] = 0;
                        prog += buf;
                    }
                    fprintf(stderr, "dcd search:   program at launch position %u: %.2f ms, list_units %llu, tmp_cap %llu, hits %u: %s\n", idx[t],
                            ns[idx[t]] * 1e-6, (unsigned long long)pl.list_units, (unsigned long long)pl.tmp_cap, w_nhits[k0 + idx[t]], prog.c_str());
                }
            }
        }
        t_phase = now();
        std::vector<u64> whoff(nw + 1, 0), wpoff(nw + 1, 0);
        for (size_t k = 0; k < nw; k++) {
            const u32 pl = order[k0 + k];
            whoff[k + 1] = whoff[k] + w_nhits[k0 + k];
            wpoff[k + 1] = wpoff[k] + w_npages[k0 + k];
            h_nhits[pl] = w_nhits[k0 + k];
            h_npages[pl] = w_npages[k0 + k];
            hoff[pl] = whoff[k];
            poff[pl] = wpoff[k];
            plan_wave[pl] = (u32)wave_host.size();   // the wave's buffers are appended below
        }
        const u64 nh = whoff[nw], npg = wpoff[nw];
        if (sink) {   // pages stay on the device: one pack kernel per wave, nothing but the counts crosses the bus
            DevBuf<u64> d_poff(pool, nw + 1);
            DevBuf<int32_t> pairs(pool, 2 * npg + 2);
            DCD_CUDA(cudaMemcpyAsync(d_poff.p, wpoff.data(), (nw + 1) * 8, cudaMemcpyHostToDevice, st));
            launch_pack_pairs(W.qb, d_poff.p, (long long)sink->page_base, pairs.p, st);
            DCD_CUDA(cudaStreamSynchronize(st));
            DCD_CUDA(cudaGetLastError());
            for (size_t k = 0; k < nw; k++) plan_pairs[order[k0 + k]] = (u64)(uintptr_t)(pairs.p + 2 * wpoff[k]);
            sink->wave_pairs.push_back(std::move(pairs));
            hits_total += nh;
            pages_total += npg;
            q_launches += 2;
            ms_download += since(t_phase);
            q_bytes += 8 * npg;
            for (u32 id : W.codes) q_bytes += 2 * (c->ix.h_code_off[id + 1] - c->ix.h_code_off[id]);
            return;
        }
        DevBuf<u64> d_hoff(pool, nw + 1), d_poff(pool, nw + 1);
        DevBuf<u32> o_page(pool, nh), o_first(pool, npg);
        DevBuf<int> o_pos(pool, nh);
        DevBuf<float> o_rank(pool, npg);
        DCD_CUDA(cudaMemcpyAsync(d_hoff.p, whoff.data(), (nw + 1) * 8, cudaMemcpyHostToDevice, st));
        DCD_CUDA(cudaMemcpyAsync(d_poff.p, wpoff.data(), (nw + 1) * 8, cudaMemcpyHostToDevice, st));
        launch_pack_results(W.qb, d_hoff.p, d_poff.p, o_page.p, o_pos.p, o_first.p, o_rank.p, st);
        wave_host.emplace_back(new WaveHost);
        WaveHost& WH = *wave_host.back();
        WH.hit_page.resize_uninit(nh);
        WH.hit_pos.resize_uninit(nh);
        WH.page_first.resize_uninit(npg);
        WH.page_rank.resize_uninit(npg);
        hits_total += nh;
        pages_total += npg;
        if (nh) {
            DCD_CUDA(cudaMemcpyAsync(WH.hit_page.data(), o_page.p, nh * 4, cudaMemcpyDeviceToHost, st));
            DCD_CUDA(cudaMemcpyAsync(WH.hit_pos.data(), o_pos.p, nh * 4, cudaMemcpyDeviceToHost, st));
        }
        if (npg) {
            DCD_CUDA(cudaMemcpyAsync(WH.page_first.data(), o_first.p, npg * 4, cudaMemcpyDeviceToHost, st));
            DCD_CUDA(cudaMemcpyAsync(WH.page_rank.data(), o_rank.p, npg * 4, cudaMemcpyDeviceToHost, st));
        }
        DCD_CUDA(cudaStreamSynchronize(st));
        DCD_CUDA(cudaGetLastError());
        q_launches += 2;
        ms_download += since(t_phase);
        if (nworkers) {   // requests whose programs are all downloaded go to the shaping threads
            std::vector<size_t> now_ready;
            for (size_t k = 0; k < nw; k++) {
                if (W.h_incomplete[k]) continue;
                const u32 qi = plan_req[order[k0 + k]];
                if (--pending[qi] == 0) now_ready.push_back(qi);
            }
            if (!now_ready.empty()) {
                { std::lock_guard<std::mutex> lk(rq_mu); ready.insert(ready.end(), now_ready.begin(), now_ready.end()); }
                rq_cv.notify_all();
            }
        }
        // algorithmic bytes, lower bound: every term list read once (2 B per ushort) + 12 B per final hit
        q_bytes += 12 * nh;
        if (W.pages_pass) return;
        for (size_t k = 0; k < nw; k++) {
            if (W.h_incomplete[k]) continue;   // counted when the program completes
            for (u32 id : plans[order[k0 + k]].codes) q_bytes += 2 * (c->ix.h_code_off[id + 1] - c->ix.h_code_off[id]);
        }
    };
    // Programs that did not reach the page cap inside their trusted prefix (finish_wave: `retry`) run again one level up; they
    // go in right behind the launch position, so they overlap the rest of the batch instead of forming its tail.
    size_t next = 0;
    size_t reruns = 0;
    for (; next < order.size() || !inflight.empty();) {
        while (next < order.size() && inflight.size() < (size_t)depth) {
            size_t k1 = next;
            u64 acc = 0;
            // a wave is cut by the scratch budget or at 512 programs: the many small programs at the end of the launch order
            // then finish (and are shaped by the host threads) in several steps instead of one long tail
            while (k1 < order.size() && (k1 == next || (acc + plan_bytes(plans[order[k1]]) <= budget && k1 - next < max_wave_programs)))
                acc += plan_bytes(plans[order[k1++]]);
            inflight.push_back(launch_wave(next, k1));
            next = k1;
        }
        finish_wave(*inflight.front());
        inflight.pop_front();
        if (!retry.empty()) {
            for (u32 pi : retry) set_level(plans[pi], plans[pi].level + 1);
            std::stable_sort(retry.begin(), retry.end(), [&](u32 a, u32 b) {
                return plans[a].work > plans[b].work;
            });
            order.insert(order.begin() + next, retry.begin(), retry.end());
            reruns += retry.size();
            retry.clear();
            w_nhits.resize(order.size() + 1, 0);
            w_npages.resize(order.size() + 1, 0);
        }
    }
    if (trace && reruns) fprintf(stderr, "dcd search: %zu program runs were repeated with a larger limit\n", reruns);
    if (ex.phase_b) {
        // ---- phase C: the hits every rank found left of its own range go to the ranks holding those pages ---------------------------
        const int R = ex.R, me = ex.me;
        auto tx = now();
        auto lap = [&](const char* what) {
            if (trace) { DCD_CUDA(cudaStreamSynchronize(s)); fprintf(stderr, "dcd search (exact): %s %.1f ms\n", what, since(tx)); tx = now(); }
        };
        for (auto x : ws.st) DCD_CUDA(cudaStreamSynchronize(x));
        std::vector<u64> foff(np + 1, 0);
        for (size_t pi = 0; pi < np; pi++) foff[pi + 1] = foff[pi] + ex.foreign_n[pi];
        DevBuf<u64> my_f(pool, foff[np] + 1);
        {
            std::vector<u64> ssrc, sdst;
            std::vector<u32> scnt;
            for (size_t pi = 0; pi < np; pi++)
                if (ex.foreign_n[pi]) { ssrc.push_back(ex.foreign_ptr[pi] / 8); sdst.push_back(foff[pi]); scnt.push_back(ex.foreign_n[pi]); }
            copy_segments(pool, nullptr, my_f.p, ssrc, sdst, scnt, s);   // absolute source addresses (in elements)
        }
        ex.wave_foreign.clear();
        DevBuf<u32> d_fn(pool, np + 1), all_fn(pool, (np + 1) * R);
        ex.foreign_n[np] = 0;
        DCD_CUDA(cudaMemcpyAsync(d_fn.p, ex.foreign_n.data(), (np + 1) * 4, cudaMemcpyHostToDevice, s));
        comm->allgather(d_fn.p, all_fn.p, (np + 1) * 4, s);
        std::vector<u32> afn((np + 1) * R);
        DCD_CUDA(cudaMemcpyAsync(afn.data(), all_fn.p, afn.size() * 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        std::vector<size_t> vcnt(R), voff(R);
        std::vector<std::vector<u64>> rf_off(R, std::vector<u64>(np + 1, 0));
        size_t vtot = 0;
        for (int r = 0; r < R; r++) {
            for (size_t pi = 0; pi < np; pi++) rf_off[r][pi + 1] = rf_off[r][pi] + afn[(size_t)r * (np + 1) + pi];
            vcnt[r] = (size_t)rf_off[r][np] * 8;
            voff[r] = vtot;
            vtot += (vcnt[r] + 15) & ~size_t(15);
        }
        DevBuf<u8> all_f(pool, vtot + 16);
        comm->allgatherv(my_f.p, all_f.p, vcnt.data(), voff.data(), s);
        DCD_CUDA(cudaStreamSynchronize(s));
        lap("foreign-hit exchange");
        // ---- phase D: own kept elements + the later ranks' elements inside this shard's range -> pages ----------------------------------
        ex.phase_b = true;
        std::vector<u64> cap(np + 1, 0);
        for (size_t pi = 0; pi < np; pi++) {
            cap[pi] = ex.own_n[pi];
            for (int r = me + 1; r < R; r++) cap[pi] += afn[(size_t)r * (np + 1) + pi];
        }
        // rank-level pointer tables (device)
        std::vector<u64> h_fvals(R), h_foffp(R), h_fcntp(R);
        std::vector<DevBuf<u64>> d_roff;
        std::vector<DevBuf<u32>> d_rcnt;
        auto launch_pages_wave = [&](size_t k0, size_t k1) {
            std::unique_ptr<WaveDev> wp(new WaveDev);
            WaveDev& W = *wp;
            const size_t nw = k1 - k0;
            W.k0 = k0;
            W.nw = nw;
            W.pages_pass = true;
            W.st = ws.st[waves_launched % ws.st.size()];
            cudaStream_t st = W.st;
            W.desc.resize(nw);
            std::vector<u64> optr(nw);
            std::vector<u32> on(nw);
            u64 hit_total = 0;
            // per rank, per wave program: offset / count of the program's foreign elements
            std::vector<u64> roff((size_t)R * nw, 0);
            std::vector<u32> rcnt((size_t)R * nw, 0);
            for (size_t k = 0; k < nw; k++) {
                const u32 pl = order[k0 + k];
                QDesc& d = W.desc[k];
                d.instr_begin = d.instr_end = 0;
                d.list_off = hit_total;
                d.tmp_off = 0;
                d.tmp_cap = 0;
                d.hit_off = hit_total;
                d.hit_cap = cap[pl];
                hit_total += cap[pl] + 1;
                optr[k] = ex.own_ptr[pl];
                on[k] = ex.own_n[pl];
                for (int r = me + 1; r < R; r++) {
                    roff[(size_t)r * nw + k] = rf_off[r][pl];
                    rcnt[(size_t)r * nw + k] = afn[(size_t)r * (np + 1) + pl];
                }
                W.codes.clear();
            }
            W.d_desc = DevBuf<QDesc>(pool, nw);
            W.d_lists = DevBuf<u64>(pool, hit_total);
            W.d_hit_page = DevBuf<u32>(pool, hit_total);
            W.d_page_first = DevBuf<u32>(pool, hit_total);
            W.d_nhits = DevBuf<u32>(pool, nw);
            W.d_npages = DevBuf<u32>(pool, nw);
            W.d_err = DevBuf<u32>(pool, 16);
            W.d_hit_pos = DevBuf<int>(pool, hit_total);
            W.d_page_rank = DevBuf<float>(pool, hit_total);
            W.d_cta_ns = DevBuf<u64>(pool, trace ? nw : 0);
            W.d_fin_ptr = DevBuf<u64>(pool, nw);          // own list pointers
            W.d_fin_n = DevBuf<u32>(pool, nw);            // own counts
            W.d_keep = DevBuf<u64>(pool, (size_t)R * nw + 3 * R);      // rank tables: offsets, then the three pointer tables
            W.d_lbase = DevBuf<u32>(pool, (size_t)R * nw);
            DCD_CUDA(cudaMemcpyAsync(W.d_desc.p, W.desc.data(), nw * sizeof(QDesc), cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemcpyAsync(W.d_fin_ptr.p, optr.data(), nw * 8, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemcpyAsync(W.d_fin_n.p, on.data(), nw * 4, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemcpyAsync(W.d_keep.p, roff.data(), roff.size() * 8, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemcpyAsync(W.d_lbase.p, rcnt.data(), rcnt.size() * 4, cudaMemcpyHostToDevice, st));
            std::vector<u64> tabs((size_t)3 * R);
            for (int r = 0; r < R; r++) {
                tabs[r] = (u64)(uintptr_t)(all_f.p + voff[r]);
                tabs[R + r] = (u64)(uintptr_t)(W.d_keep.p + (size_t)r * nw);
                tabs[2 * R + r] = (u64)(uintptr_t)(W.d_lbase.p + (size_t)r * nw);
            }
            u64* d_tabs = W.d_keep.p + (size_t)R * nw;
            DCD_CUDA(cudaMemcpyAsync(d_tabs, tabs.data(), tabs.size() * 8, cudaMemcpyHostToDevice, st));
            DCD_CUDA(cudaMemsetAsync(W.d_err.p, 0, 64, st));
            DCD_CUDA(cudaStreamSynchronize(st));   // host vectors above
            QBatch& qb = W.qb;
            qb.desc = W.d_desc.p;
            qb.nqueries = (u32)nw;
            qb.page_end = c->ix.d_page_end.p;
            qb.npages = c->ix.npages;
            qb.coord_origin = c->ix.coord_origin;
            qb.max_pages = (u32)std::max(0, c->opt.max_found_pages);
            qb.hit_page = W.d_hit_page.p;
            qb.hit_pos = W.d_hit_pos.p;
            qb.page_first = W.d_page_first.p;
            qb.page_rank = W.d_page_rank.p;
            qb.nhits = W.d_nhits.p;
            qb.npages_found = W.d_npages.p;
            qb.error = W.d_err.p;
            PagesPass pp{};
            pp.own_ptr = reinterpret_cast<const u64* const*>(W.d_fin_ptr.p);
            pp.own_n = W.d_fin_n.p;
            pp.nranks = R;
            pp.rank = me;
            pp.f_vals = reinterpret_cast<const u64* const*>(d_tabs);
            pp.f_off = reinterpret_cast<const u64* const*>(d_tabs + R);
            pp.f_cnt = reinterpret_cast<const u32* const*>(d_tabs + 2 * R);
            pp.range_end = ex.range_end;
            pp.work = W.d_lists.p;
            launch_query_pages(qb, pp, st);
            DCD_CUDA(cudaEventCreate(&W.ev1));
            DCD_CUDA(cudaEventRecord(W.ev1, st));
            waves_launched++;
            return wp;
        };
        u64 d_budget = std::max<u64>(budget, 64ull << 20);
        for (size_t k0 = 0; k0 < np;) {
            size_t k1 = k0;
            u64 acc = 0;
            while (k1 < np && (k1 == k0 || (acc + (cap[order[k1]] + 1) * 24 <= d_budget && k1 - k0 < 4096))) acc += (cap[order[k1++]] + 1) * 24;
            auto wp = launch_pages_wave(k0, k1);
            finish_wave(*wp);
            k0 = k1;
        }
        lap("pages pass");
        ex.wave_own.clear();
    }
    if (np) {
        DCD_CUDA(cudaMemcpyAsync(h_nfound.data(), d_nfound.p, (nslots + 1) * 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        for (const Plan& pl : plans) {   // limited programs: wordInfo.nFound from the per-code value counts
            if (pl.base_instr.empty()) continue;
            for (size_t i = 1; i < pl.instr.size(); i++)
                if (pl.instr[i].op == QOP_COUNT && (pl.instr[i - 1].op == QOP_LOAD || pl.instr[i - 1].op == QOP_LOADC))
                    h_nfound[pl.instr[i].a] = pl.instr[i - 1].b ? c->ix.h_code_nvals[pl.codes[pl.instr[i - 1].a]] : 0u;
        }
        c->counters.kernel_launches = q_launches + q_launches_pre;
        c->counters.search_ms_device = q_ms;
        c->counters.search_hits = hits_total;
        c->counters.search_bytes = q_bytes;
    }

    if (sink) {
        sink->counts.assign(nq, 0);
        sink->req_src.assign(nq, 0);
        sink->total = 0;
        for (size_t qi = 0; qi < nq; qi++) {
            out->q[qi].status = prep[qi].status;
            const int plan = prep[qi].main_plan >= 0 ? prep[qi].main_plan : prep[qi].field_plan;   // Search.cs:538
            if (prep[qi].status || plan < 0) continue;
            sink->counts[qi] = (int32_t)h_npages[plan];
            sink->req_src[qi] = plan_pairs[plan];
            sink->total += h_npages[plan];
        }
        join_workers.close();
        if (trace)
            fprintf(stderr, "dcd search (pages on device): %zu requests, %zu programs: plan %.1f ms, wave setup %.1f ms, kernel %.1f ms, pack %.1f ms\n",
                    nq, np, ms_plan, ms_wave_prep, ms_kernel_wait, ms_download);
        return;
    }
    t_phase = now();
    // whatever the worker threads have not shaped (general path, small batches) is shaped here; then the per-term counts
    join_workers.close();
    for (auto& e : worker_err) if (e) std::rethrow_exception(e);
    for (size_t qi = 0; qi < nq; qi++) {
        if (!nworkers || prep[qi].status || pending[qi] != 0 || (prep[qi].main_plan < 0 && prep[qi].field_plan < 0)) finish_query(qi);
        if (!prep[qi].status && (prep[qi].main_plan >= 0 || prep[qi].field_plan >= 0))
            for (auto& t : prep[qi].terms) out->q[qi].words.push_back({t.word, (int)h_nfound[t.nfound_slot]});
    }
    ms_shape = since(t_phase);
    if (trace) {
        u64 tot_units = 0;
        std::vector<u64> sz;
        for (auto& pl : plans) { tot_units += pl.list_units; sz.push_back(pl.list_units); }
        std::sort(sz.rbegin(), sz.rend());
        fprintf(stderr, "dcd search: list units total %llu, largest programs:", (unsigned long long)tot_units);
        for (size_t i = 0; i < std::min<size_t>(sz.size(), 8); i++) fprintf(stderr, " %llu", (unsigned long long)sz[i]);
        fprintf(stderr, "\n");
    }
    if (trace)
        fprintf(stderr, "dcd search: %zu requests, %zu programs: plan %.1f ms, wave setup %.1f ms, kernel %.1f ms, pack+download %.1f ms, "
                        "shape %.1f ms\n", nq, np, ms_plan, ms_wave_prep, ms_kernel_wait, ms_download, ms_shape);
}

void engine_pages_fetch(dcd_ctx_impl* c, PagesSink& sink, int32_t* dst, u64 cap_pairs, bool on_device) {
    if (sink.total > cap_pairs) throw StatusError(DCD_E_INVALID, "output buffer too small");
    cudaStream_t s = c->stream;
    DevPool& pool = c->pool;
    const size_t nq = sink.counts.size();
    if (sink.total) {
        std::vector<u64> off(nq);
        std::vector<u32> cnt(nq);
        u64 at = 0;
        for (size_t i = 0; i < nq; i++) { off[i] = at; cnt[i] = (u32)sink.counts[i]; at += cnt[i]; }
        DevBuf<u64> d_src(pool, nq), d_off(pool, nq);
        DevBuf<u32> d_cnt(pool, nq);
        DCD_CUDA(cudaMemcpyAsync(d_src.p, sink.req_src.data(), nq * 8, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(d_off.p, off.data(), nq * 8, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(d_cnt.p, cnt.data(), nq * 4, cudaMemcpyHostToDevice, s));
        if (on_device) {
            launch_gather_pairs(d_src.p, d_off.p, d_cnt.p, (u32)nq, dst, s);
        } else {
            DevBuf<int32_t> tmp(pool, 2 * sink.total);
            launch_gather_pairs(d_src.p, d_off.p, d_cnt.p, (u32)nq, tmp.p, s);
            DCD_CUDA(cudaMemcpyAsync(dst, tmp.p, sink.total * 8, cudaMemcpyDeviceToHost, s));
        }
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
    }
    sink.wave_pairs.clear();
    sink.req_src.clear();
}

}  // namespace dcd
