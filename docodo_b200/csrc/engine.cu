// Host orchestration of the device build pipeline (Builder.Save + MergeIndexes + Load of the reference:
// Build.cs:370-434, Index.cs:488-512, 695-752) and the ordinal ranking of the code strings.
#include "engine.h"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "encode.cuh"
#include "shard.h"
#include "sort.cuh"

namespace dcd {

// ---------------------------------------------------------------------------------------------------
// memory pool
// ---------------------------------------------------------------------------------------------------
void DevPool::add_free(char* p) { free_by_size.emplace(segs[p].size, p); }
void DevPool::remove_free(char* p) {
    auto r = free_by_size.equal_range(segs[p].size);
    for (auto it = r.first; it != r.second; ++it)
        if (it->second == p) { free_by_size.erase(it); return; }
}
void* DevPool::alloc(size_t bytes) {
    bytes = (bytes + 511) & ~size_t(511);
    auto it = free_by_size.lower_bound(bytes);
    if (it != free_by_size.end()) {
        char* p = it->second;
        free_by_size.erase(it);
        Seg& sg = segs[p];
        sg.free = false;
        bytes_cached -= sg.size;
        if (sg.size > bytes + bytes / 8 + (1 << 20)) {   // split: the tail stays free
            char* q = p + bytes;
            segs[q] = Seg{sg.size - bytes, true, sg.slab};
            sg.size = bytes;
            bytes_cached += segs[q].size;
            add_free(q);
        }
        bytes_live += sg.size;
        peak = std::max(peak, bytes_live);
        return p;
    }
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        trim();
        e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            cudaGetLastError();
            throw StatusError(DCD_E_CUDA, "cudaMalloc of " + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(e));
        }
    }
    slabs[(char*)p] = bytes;
    segs[(char*)p] = Seg{bytes, false, (char*)p};
    bytes_live += bytes;
    peak = std::max(peak, bytes_live);
    return p;
}
void DevPool::release(void* vp) {
    char* p = (char*)vp;
    auto it = segs.find(p);
    if (it == segs.end() || it->second.free) return;
    it->second.free = true;
    bytes_live -= it->second.size;
    bytes_cached += it->second.size;
    // coalesce with the next and the previous segment of the same slab
    auto nx = std::next(it);
    if (nx != segs.end() && nx->second.free && nx->second.slab == it->second.slab && nx->first == p + it->second.size) {
        remove_free(nx->first);
        it->second.size += nx->second.size;
        segs.erase(nx);
    }
    if (it != segs.begin()) {
        auto pv = std::prev(it);
        if (pv->second.free && pv->second.slab == it->second.slab && pv->first + pv->second.size == p) {
            remove_free(pv->first);
            pv->second.size += it->second.size;
            segs.erase(it);
            it = pv;
        }
    }
    add_free(it->first);
}
void DevPool::shrink(void* vp, size_t new_bytes) {
    char* p = (char*)vp;
    auto it = segs.find(p);
    if (it == segs.end() || it->second.free) return;
    new_bytes = (new_bytes + 511) & ~size_t(511);
    if (new_bytes + (1 << 20) >= it->second.size) return;
    char* q = p + new_bytes;
    const size_t tail = it->second.size - new_bytes;
    it->second.size = new_bytes;
    segs[q] = Seg{tail, false, it->second.slab};   // a live segment, released through the common path (coalescing)
    release(q);
}
void DevPool::trim() {
    for (auto sl = slabs.begin(); sl != slabs.end();) {
        auto it = segs.find(sl->first);
        if (it != segs.end() && it->second.free && it->second.size == sl->second) {
            remove_free(it->first);
            bytes_cached -= it->second.size;
            segs.erase(it);
            cudaFree(sl->first);
            sl = slabs.erase(sl);
        } else {
            ++sl;
        }
    }
}
DevPool::~DevPool() {
    for (auto& kv : slabs) cudaFree(kv.first);
}

// ---------------------------------------------------------------------------------------------------
// pinned host block cache
// ---------------------------------------------------------------------------------------------------
static std::multimap<size_t, void*>& pinned_cache() { static std::multimap<size_t, void*> c; return c; }
static std::mutex& pinned_mutex() { static std::mutex m; return m; }
void* pinned_cache_alloc(size_t bytes, size_t* cap_out) {
    bytes = (bytes + 4095) & ~size_t(4095);
    {
        std::lock_guard<std::mutex> g(pinned_mutex());
        auto it = pinned_cache().lower_bound(bytes);
        if (it != pinned_cache().end() && it->first <= 2 * bytes + (1 << 20)) {
            void* p = it->second;
            *cap_out = it->first;
            pinned_cache().erase(it);
            return p;
        }
    }
    void* p = nullptr;
    cudaError_t e = cudaMallocHost(&p, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); throw StatusError(DCD_E_CUDA, std::string("cudaMallocHost failed: ") + cudaGetErrorString(e)); }
    *cap_out = bytes;
    return p;
}
void pinned_cache_release(void* p, size_t cap) {
    std::lock_guard<std::mutex> g(pinned_mutex());
    pinned_cache().emplace(cap, p);
}

void StrTableOwner::create(DevPool& pool, u32 cap_log2, u32 max_entries, u32 arena_units, cudaStream_t s) {
    const size_t cap = size_t(1) << cap_log2;
    slot_hash = DevBuf<u64>(pool, cap);
    slot_entry = DevBuf<u32>(pool, cap);
    ent_off = DevBuf<u32>(pool, max_entries);
    ent_len = DevBuf<u32>(pool, max_entries);
    ent_flags = DevBuf<u32>(pool, max_entries);
    ent_value = DevBuf<u32>(pool, max_entries);
    arena = DevBuf<u16>(pool, arena_units);
    counters = DevBuf<u32>(pool, 8);
    view.slot_hash = slot_hash.p;
    view.slot_entry = slot_entry.p;
    view.ent_off = ent_off.p;
    view.ent_len = ent_len.p;
    view.ent_flags = ent_flags.p;
    view.ent_value = ent_value.p;
    view.arena = arena.p;
    view.counters = counters.p;
    view.cap_mask = (u32)(cap - 1);
    view.max_entries = max_entries;
    view.arena_cap = arena_units;
    clear(s);
}
void StrTableOwner::clear(cudaStream_t s) {
    DCD_CUDA(cudaMemsetAsync(slot_hash.p, 0, slot_hash.bytes(), s));
    DCD_CUDA(cudaMemsetAsync(slot_entry.p, 0xFF, slot_entry.bytes(), s));
    DCD_CUDA(cudaMemsetAsync(counters.p, 0, counters.bytes(), s));
    DCD_CUDA(cudaMemsetAsync(arena.p, 0, arena.bytes(), s));
    DCD_CUDA(cudaMemsetAsync(ent_len.p, 0, ent_len.bytes(), s));
}

// ---------------------------------------------------------------------------------------------------
// small kernels of the ranking / finishing stages
// ---------------------------------------------------------------------------------------------------
// 4 UTF-16 units [4c, 4c+4) of entry perm[i], big-endian packed (missing units = 0 so that a prefix sorts first:
// String.CompareOrdinal, Index.cs:142-149)
__global__ void k_chunk_keys(StrTable t, const u32* __restrict__ perm, u32 n, int chunk, u64* __restrict__ keys) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 e = perm[i];
    const u32 len = t.ent_len[e];
    const u16* s = t.arena + t.ent_off[e];
    u64 k = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const u32 p = (u32)chunk * 4 + j;
        k = (k << 16) | (p < len ? (u64)s[p] : 0ull);
    }
    keys[i] = k;
}
__global__ void k_max_len(StrTable t, const u32* __restrict__ perm, u32 n, u32* __restrict__ out) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    u32 l = i < n ? t.ent_len[perm[i]] : 0;
    for (int o = 16; o > 0; o >>= 1) l = max(l, __shfl_xor_sync(0xffffffffu, l, o));
    if (lane_id() == 0 && l) atomicMax(out, l);
}
__global__ void k_assign_ranks(StrTable t, const u32* __restrict__ perm, u32 n, u64* __restrict__ str_len) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 e = perm[i];
    t.ent_value[e] = i;
    str_len[i] = t.ent_len[e];
}
// code strings in rank order, packed: off_incl[i] = inclusive prefix of lengths
__global__ void k_pack_codes(StrTable t, const u32* __restrict__ perm, u32 n, const u64* __restrict__ off_incl,
                             u16* __restrict__ chars, u32* __restrict__ off32) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 e = perm[i];
    const u32 len = t.ent_len[e];
    const u64 end = off_incl[i];
    const u16* s = t.arena + t.ent_off[e];
    u16* d = chars + (end - len);
    for (u32 j = 0; j < len; j++) d[j] = s[j];
    off32[i + 1] = (u32)end;
    if (i == 0) off32[0] = 0;
}
// sharded build: dense, order-preserving LOCAL code ids (rank among the codes present on this rank).  The posting sort runs
// on the local ids (fewer key bits than the global dictionary needs); glob[] maps them back when the lists are laid out.
__global__ void k_or_flag(StrTable t, const u32* __restrict__ perm, u32 n, u32 flag) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) t.ent_flags[perm[i]] |= flag;
}
__global__ void k_present(StrTable t, const u32* __restrict__ perm, u32 n, u64* __restrict__ present) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) present[i] = (t.ent_flags[perm[i]] & EF_LOCAL) ? 1 : 0;
}
__global__ void k_local_ids(StrTable t, const u32* __restrict__ perm, u32 n, const u64* __restrict__ present_incl, u32* __restrict__ glob,
                            u32* __restrict__ local_of_glob) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 e = perm[i];
    if (t.ent_flags[e] & EF_LOCAL) {
        const u32 l = (u32)present_incl[i] - 1;
        glob[l] = i;
        local_of_glob[i] = l;
        t.ent_value[e] = l;
    } else {
        local_of_glob[i] = 0xFFFFFFFFu;
    }
}
__global__ void k_base_by_local(const u64* __restrict__ base, const u32* __restrict__ glob, u32 n, u64* __restrict__ out) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = base[glob[i]];
}
__global__ void k_iota(u32* p, u32 n) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}
__global__ void k_keep_flags(const u64* __restrict__ adv, u64 n, u64* __restrict__ keep) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keep[i] = adv[i] > 0 ? 1 : 0;
}
__global__ void k_compact_pages(const u64* __restrict__ adv, const u64* __restrict__ keep_incl, const u64* __restrict__ coord_end,
                                u64 n, u64* __restrict__ out_end, u32* __restrict__ out_orig) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || adv[i] == 0) return;
    const u64 k = keep_incl[i] - 1;
    out_end[k] = coord_end[i];
    if (out_orig) out_orig[k] = (u32)i;
}
// out[i] = ushorts in front of key tile want[i] (prefix[] = inclusive per-tile prefixes of the lengths pass)
__global__ void k_gather_prefix(const u64* __restrict__ prefix, const u64* __restrict__ want, u64 n, u64* __restrict__ out) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = want[i] ? (prefix[want[i] - 1] & LB_VALUE_MASK) : 0;
}
__global__ void k_max_key(const u64* __restrict__ keys, u64 n, u64* __restrict__ out) {
    // Builder.maxCoord == coordinate of the last posting added; postings are emitted in coordinate order, so that
    // is the maximum coordinate over all keys (body text) -- reduced here over the unsorted tail-free key array.
    u64 m = 0;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) m = max(m, keys[i] & COORD_MASK);
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane_id() == 0) atomicMax(out, m);
}

static inline unsigned grid_for(u64 n, int block = 256) { return (unsigned)((n + block - 1) / block); }
void launch_flag_extras(StrTable t, const u32* entries, const u8* kinds, u32 n, cudaStream_t s);

static int bits_for(u64 n) {   // bits needed to represent ids 0..n-1
    int b = 1;
    while (b < 63 && (1ull << b) < n) b++;
    return b;
}

struct StageTimer {
    cudaStream_t s;
    cudaEvent_t ev[16];
    int n = 0;
    explicit StageTimer(cudaStream_t st) : s(st) { for (auto& e : ev) cudaEventCreate(&e); }
    ~StageTimer() { for (auto& e : ev) cudaEventDestroy(e); }
    void mark() { if (n < 16) cudaEventRecord(ev[n++], s); }
    float ms(int a, int b) { float f = 0; cudaEventElapsedTime(&f, ev[a], ev[b]); return f; }
};

// ---------------------------------------------------------------------------------------------------
// build
// ---------------------------------------------------------------------------------------------------
void engine_build_finish(dcd_build_impl* b, ShardComm* comm) {
    const auto wall0 = std::chrono::steady_clock::now();
    dcd_ctx_impl* c = b->ctx;
    cudaStream_t s = c->stream;
    DevPool& pool = c->pool;
    DCD_CUDA(cudaSetDevice(c->device));
    c->ix.clear();
    c->merged.clear();
    c->counters.shard_p2p = 0;
    u64 launches = 0;
    StageTimer tm(s);
    StageTimer xt(s);   // sharded build: [0,1] dictionary unification, [2,3] first-delta bases, [4,5] segment exchange + concat
    if (comm) comm->bytes_moved = 0;
    tm.mark();   // 0

    // ---- code table (stop words + every code string) ----------------------------------------------
    const u32 cap_log2 = c->opt.dict_capacity_log2 ? c->opt.dict_capacity_log2 : 23;
    const u32 max_entries = (u32)((size_t(1) << cap_log2) / 2);
    const u32 arena_units = max_entries * 12u;
    StrTableOwner tab;
    tab.create(pool, cap_log2, max_entries, arena_units, s);

    BuildParams bp{};
    bp.tab = tab.view;
    bp.nvocs = (int)c->vocs.size();
    for (int v = 0; v < bp.nvocs; v++) {
        bp.vocs[v].tab = c->vocs[v]->dev.view;
        bp.vocs[v].begin = (u16)c->vocs[v]->begin;
        bp.vocs[v].end = (u16)c->vocs[v]->end;
        bp.vocs[v].stemmer = c->vocs[v]->stemmer;
    }
    bp.stemmers_mask = c->opt.stemmers_mask;

    auto upload_strings = [&](const std::vector<u16>& chars, const std::vector<u32>& offs, DevBuf<u16>& dch, DevBuf<u32>& dof) {
        dch = DevBuf<u16>(pool, chars.size());
        dof = DevBuf<u32>(pool, offs.size());
        if (!chars.empty()) DCD_CUDA(cudaMemcpyAsync(dch.p, chars.data(), chars.size() * 2, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(dof.p, offs.data(), offs.size() * 4, cudaMemcpyHostToDevice, s));
    };
    // stop words (Index.stopWords, Index.cs:226-235) -> EF_STOP entries
    {
        std::vector<u16> chars;
        std::vector<u32> offs{0};
        for (const auto& w : c->stop) {
            if (w.empty() || w.size() > (size_t)MAX_KEY_UNITS) continue;
            chars.insert(chars.end(), (const u16*)w.data(), (const u16*)w.data() + w.size());
            offs.push_back((u32)chars.size());
        }
        if (offs.size() > 1) {
            DevBuf<u16> dch;
            DevBuf<u32> dof;
            upload_strings(chars, offs, dch, dof);
            launch_insert_strings(tab.view, dch.p, dof.p, (u32)offs.size() - 1, EF_STOP, nullptr, nullptr, s);
            launches++;
            DCD_CUDA(cudaStreamSynchronize(s));
        }
    }

    // ---- K1/K2: one pass over the text -> token records + distinct-word table -----------------------
    DevBuf<u32> ticket(pool, 16);
    DevBuf<u64> counters64(pool, 16);   // [0] postings, [2] tokens, [3] max coord, [4] arena total
    DevBuf<u32> d_err(pool, 16);
    const u32 tok_cap = (u32)(TK_TILE / std::max(1, c->opt.min_word_length) + 1);
    u64 text_units = 0;
    struct BatchTok { DevBuf<u64> tok; DevBuf<u32> tile_count, page_bits, rec; DevBuf<u16> cnt16; DevBuf<i64> tile_delta; bool v2 = false; };
    std::vector<BatchTok> btok(b->batches.size());
    bool any_v2 = false;
    for (size_t bi = 0; bi < b->batches.size(); bi++) {
        TextBatch& bt = b->batches[bi];
        bt.ntiles = (u32)((bt.n_units + TK_TILE - 1) / TK_TILE);
        text_units += bt.n_units;
        if (!bt.ntiles) continue;
        if (bt.uniform && c->opt.min_word_length >= 3 && c->opt.max_word_length <= 32) {
            // bulk pages: warp-autonomous tokenizer, 4-byte records (tokenize.cuh, v2)
            btok[bi].v2 = any_v2 = true;
            bt.ntiles2 = (u32)((bt.n_units + T2_OWN - 1) / T2_OWN);
            btok[bi].cnt16 = DevBuf<u16>(pool, bt.ntiles2);
            btok[bi].rec = DevBuf<u32>(pool, (size_t)bt.ntiles2 * T2_CAP);
            btok[bi].page_bits = DevBuf<u32>(pool, (size_t)(bt.n_units / 32) + 72);
            DCD_CUDA(cudaMemsetAsync(btok[bi].page_bits.p, 0, btok[bi].page_bits.bytes(), s));
            launch_page_bits(bt.page_off, bt.npages, btok[bi].page_bits.p, s);
            launches += 1;
            continue;
        }
        bt.tile_first_page = DevBuf<u32>(pool, bt.ntiles);
        btok[bi].tile_delta = DevBuf<i64>(pool, bt.ntiles);
        btok[bi].tile_count = DevBuf<u32>(pool, bt.ntiles);
        btok[bi].tok = DevBuf<u64>(pool, (size_t)bt.ntiles * tok_cap);
        btok[bi].page_bits = DevBuf<u32>(pool, (size_t)bt.ntiles * (TK_TILE / 32) + 8);
        DCD_CUDA(cudaMemsetAsync(btok[bi].page_bits.p, 0, btok[bi].page_bits.bytes(), s));
        launch_tile_info(bt.page_off, bt.page_delta.p, bt.npages, bt.n_units, bt.tile_first_page.p, btok[bi].tile_delta.p, bt.ntiles, s);
        launch_page_bits(bt.page_off, bt.npages, btok[bi].page_bits.p, s);
        launches += 2;
    }
    const u32 wcap_max = any_v2 ? 22u : 23u;   // v2 records hold 22-bit slots
    u32 wcap_log2 = std::min<u32>(wcap_max, std::max<u32>(16u, cap_log2 - 1));
    u32 warena_units = 1u << 22;
    DevBuf<ulonglong2> wslots;
    DevBuf<u16> warena;
    DevBuf<u32> wcounters(pool, 8);
    WordTable wt{};
    u32 n_words = 0;
    u64 n_tokens = 0;
    for (int attempt = 0;; attempt++) {
        wslots = DevBuf<ulonglong2>(pool, size_t(1) << wcap_log2);
        warena = DevBuf<u16>(pool, warena_units);
        wt.slots = wslots.p;
        wt.cap_mask = (u32)((size_t(1) << wcap_log2) - 1);
        wt.arena = warena.p;
        wt.arena_cap = warena_units;
        wt.counters = wcounters.p;
        DCD_CUDA(cudaMemsetAsync(wslots.p, 0, wslots.bytes(), s));
        // the last slot is never a word: its index is the v2 token records' "no word" marker (T2_SLOT_NONE) when the table
        // has 2^22 slots.  An all-ones sentinel (tag 7) matches no key, so probes walk past it; k_list_words skips it.
        DCD_CUDA(cudaMemsetAsync(wslots.p + wt.cap_mask, 0xFF, sizeof(ulonglong2), s));
        DCD_CUDA(cudaMemsetAsync(wcounters.p, 0, wcounters.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(counters64.p, 0, counters64.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(d_err.p, 0, 64, s));
        for (size_t bi = 0; bi < b->batches.size(); bi++) {
            TextBatch& bt = b->batches[bi];
            if (!bt.ntiles) continue;
            if (btok[bi].v2) {
                Tok2Params t2{};
                t2.text = bt.text;
                t2.enc = bt.enc;
                t2.n_units = bt.n_units;
                t2.page_bits = btok[bi].page_bits.p;
                t2.ntiles = bt.ntiles2;
                t2.tab = wt;
                t2.cls = c->d_cls.p;
                t2.lower = c->d_lower.p;
                t2.min_len = c->opt.min_word_length;
                t2.max_len = c->opt.max_word_length;
                t2.rec = btok[bi].rec.p;
                t2.tile_count = btok[bi].cnt16.p;
                t2.token_total = counters64.p + 2;
                if (bt.h2d && !bt.h2d->ev.empty()) {
                    // host text still arriving: tokenize the tiles whose window [960 t - 1, 960 t + T2_WIN) is resident
                    u32 t_begin = 0;
                    for (size_t j = 0; j < bt.h2d->ev.size(); j++) {
                        const bool last_chunk = j + 1 == bt.h2d->ev.size();
                        const u64 ready = bt.h2d->ev[j].first;
                        u32 t_end = bt.ntiles2;
                        if (!last_chunk) t_end = ready >= (u64)T2_WIN ? (u32)std::min<u64>(bt.ntiles2, (ready - T2_WIN) / T2_OWN + 1) : 0u;
                        if (t_end <= t_begin) continue;
                        DCD_CUDA(cudaStreamWaitEvent(s, bt.h2d->ev[j].second, 0));
                        t2.tile_begin = t_begin;
                        t2.ntiles = t_end;
                        launch_tokenize2(t2, c->nsm, s);
                        launches++;
                        t_begin = t_end;
                    }
                    continue;
                }
                launch_tokenize2(t2, c->nsm, s);
                launches++;
                continue;
            }
            if (bt.h2d && !bt.h2d->ev.empty()) DCD_CUDA(cudaStreamWaitEvent(s, bt.h2d->ev.back().second, 0));
            TextView tv{bt.text, bt.enc, bt.n_units, bt.page_off, bt.page_delta.p, bt.tile_first_page.p, btok[bi].page_bits.p, btok[bi].tile_delta.p,
                        bt.npages, bt.ntiles};
            TokenizeParams tp{};
            tp.tab = wt;
            tp.cls = c->d_cls.p;
            tp.lower = c->d_lower.p;
            tp.min_len = c->opt.min_word_length;
            tp.max_len = c->opt.max_word_length;
            tp.tok = btok[bi].tok.p;
            tp.tok_cap = tok_cap;
            tp.tile_count = btok[bi].tile_count.p;
            tp.token_total = counters64.p + 2;
            tp.error = d_err.p;
            launch_tokenize(tv, tp, c->nsm, s);
            launches++;
        }
        u32 hw[8];
        u64 h64a[16];
        DCD_CUDA(cudaMemcpyAsync(hw, wcounters.p, sizeof hw, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaMemcpyAsync(h64a, counters64.p, sizeof h64a, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        n_words = hw[2];
        n_tokens = h64a[2];
        const bool full = hw[1] != 0 || n_words > (wt.cap_mask + 1) / 2;
        if (!full) break;
        if (wcap_log2 >= wcap_max && warena_units >= (1u << 28))
            throw StatusError(DCD_E_CAPACITY, "too many distinct words for the word table: split the corpus into document-range shards");
        wcap_log2 = std::min<u32>(wcap_max, wcap_log2 + 1);      // grow and tokenize again (rare: only the first build of a
        if (warena_units < (1u << 28)) warena_units <<= 2;  // much larger vocabulary pays this)
    }
    {
        u32 herr0[16];
        DCD_CUDA(cudaMemcpyAsync(herr0, d_err.p, 64, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        if (herr0[2]) throw StatusError(DCD_E_CAPACITY, "coordinate exceeds 2^40");
    }
    // host-supplied words / codes (header pages, Builder.AddWord / Builder.Add)
    const u32 n_extra = (u32)b->extra_coord.size();
    DevBuf<u32> d_extra_entries(pool, n_extra);
    DevBuf<u64> d_extra_coord(pool, n_extra);
    DevBuf<u8> d_extra_kind(pool, n_extra);
    DevBuf<u16> d_extra_chars;
    DevBuf<u32> d_extra_offs;
    if (n_extra) {
        upload_strings(b->extra_chars, b->extra_offs, d_extra_chars, d_extra_offs);
        launch_insert_strings(tab.view, d_extra_chars.p, d_extra_offs.p, n_extra, 0, nullptr, d_extra_entries.p, s);
        DCD_CUDA(cudaMemcpyAsync(d_extra_coord.p, b->extra_coord.data(), n_extra * 8, cudaMemcpyHostToDevice, s));
        DCD_CUDA(cudaMemcpyAsync(d_extra_kind.p, b->extra_kind.data(), n_extra, cudaMemcpyHostToDevice, s));
        launch_flag_extras(tab.view, d_extra_entries.p, d_extra_kind.p, n_extra, s);
        launches += 2;
        DCD_CUDA(cudaStreamSynchronize(s));
    }
    DCD_CUDA(cudaGetLastError());
    tm.mark();   // 1: tokenize done

    // ---- K4: codes per distinct word ------------------------------------------------------------------
    DevBuf<u32> word_slots(pool, n_words), d_small(pool, 16);
    DevBuf<u32> word_codes(pool, (size_t)n_words * 8), extra_codes(pool, (size_t)n_extra * 8);
    DevBuf<u64> slot_codes(pool, size_t(1) << wcap_log2);
    DCD_CUDA(cudaMemsetAsync(d_small.p, 0, 64, s));
    launch_list_words(wt, word_slots.p, d_small.p, s);
    launch_word_codes(bp, wt, word_slots.p, nullptr, nullptr, nullptr, n_words, word_codes.p, s);
    launch_word_codes(bp, wt, nullptr, d_extra_chars.p, d_extra_offs.p, d_extra_kind.p, n_extra, extra_codes.p, s);
    launches += 3;
    u32 hc[8];
    DCD_CUDA(cudaMemcpyAsync(hc, tab.counters.p, sizeof hc, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    DCD_CUDA(cudaGetLastError());
    if (hc[2]) throw StatusError(DCD_E_CAPACITY, "code table full while deriving codes: raise dcd_options.dict_capacity_log2");

    // ---- sharded build: unify the code dictionaries --------------------------------
    // Every rank packs its code strings, the packed tables are all-gathered and the strings of the other ranks are
    // inserted into the local table, so that the ordinal ranking below yields identical GLOBAL code ids everywhere
    // (what the k-way merge by ordinal key of MergeFiles does with strings, Index.cs:600-689).
    xt.mark();   // x0
    if (comm && comm->nranks > 1) {
        const int R = comm->nranks;
        const u32 n_ent0 = hc[0];
        DevBuf<u32> perm(pool, n_ent0);
        DCD_CUDA(cudaMemsetAsync(d_small.p, 0, 64, s));
        launch_gather_flagged(tab.view, n_ent0, EF_CODE, perm.p, d_small.p, s);
        u32 n_mine = 0;
        DCD_CUDA(cudaMemcpyAsync(&n_mine, d_small.p, 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        if (n_mine) k_or_flag<<<grid_for(n_mine), 256, 0, s>>>(tab.view, perm.p, n_mine, EF_LOCAL);
        DevBuf<u64> lens(pool, n_mine), lens_incl(pool, n_mine);
        DevBuf<u32> my_off(pool, (size_t)n_mine + 1);
        u64 my_chars = 0;
        DevBuf<u16> my_chars_buf;
        if (n_mine) {
            k_assign_ranks<<<grid_for(n_mine), 256, 0, s>>>(tab.view, perm.p, n_mine, lens.p);
            DevBuf<u64> lb(pool, (n_mine + 2047) / 2048 + 1);
            DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
            DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
            launch_inclusive_scan_u64(lens.p, lens_incl.p, n_mine, 0, lb.p, ticket.p, s);
            DCD_CUDA(cudaMemcpyAsync(&my_chars, lens_incl.p + (n_mine - 1), 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            my_chars_buf = DevBuf<u16>(pool, my_chars);
            k_pack_codes<<<grid_for(n_mine), 256, 0, s>>>(tab.view, perm.p, n_mine, lens_incl.p, my_chars_buf.p, my_off.p);
            launches += 4;
        } else {
            my_chars_buf = DevBuf<u16>(pool, 1);
            DCD_CUDA(cudaMemsetAsync(my_off.p, 0, 4, s));
        }
        // sizes of every rank
        DevBuf<u64> d_meta(pool, 2), d_meta_all(pool, (size_t)2 * R);
        const u64 meta[2] = {n_mine, my_chars};
        DCD_CUDA(cudaMemcpyAsync(d_meta.p, meta, 16, cudaMemcpyHostToDevice, s));
        comm->allgather(d_meta.p, d_meta_all.p, 16, s);
        std::vector<u64> meta_all((size_t)2 * R);
        DCD_CUDA(cudaMemcpyAsync(meta_all.data(), d_meta_all.p, meta_all.size() * 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        std::vector<size_t> cn(R), co(R), on(R), oo(R);
        size_t ctot = 0, otot = 0;
        for (int r = 0; r < R; r++) {
            cn[r] = (size_t)meta_all[2 * r + 1] * 2; co[r] = ctot; ctot += (cn[r] + 15) & ~size_t(15);
            on[r] = ((size_t)meta_all[2 * r] + 1) * 4; oo[r] = otot; otot += (on[r] + 15) & ~size_t(15);
        }
        DevBuf<u8> all_chars(pool, ctot + 16), all_offs(pool, otot + 16);
        comm->allgatherv(my_chars_buf.p, all_chars.p, cn.data(), co.data(), s);
        comm->allgatherv(my_off.p, all_offs.p, on.data(), oo.data(), s);
        if (R > 16) throw StatusError(DCD_E_UNSUPPORTED, "more than 16 shards");
        InsertSources isrc{};
        for (int r = 0; r < R; r++) {
            if (r == comm->rank || !meta_all[2 * r]) continue;
            isrc.chars_off[isrc.n] = co[r];
            isrc.offs_off[isrc.n] = oo[r];
            isrc.first[isrc.n + 1] = isrc.first[isrc.n] + (u32)meta_all[2 * r];
            isrc.n++;
        }
        launch_insert_strings_multi(tab.view, all_chars.p, all_offs.p, isrc, EF_CODE, s);
        launches++;
        DCD_CUDA(cudaMemcpyAsync(hc, tab.counters.p, sizeof hc, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        if (hc[2]) throw StatusError(DCD_E_CAPACITY, "code table full while unifying the shard dictionaries: raise dcd_options.dict_capacity_log2");
    }
    xt.mark();   // x1
    const u32 n_entries = hc[0];
    const u32 n_words_entries = n_words;

    // ---- ordinal ranking of the code strings ----------------------------------------------------------
    DevBuf<u32> perm_a(pool, n_entries), perm_b(pool, n_entries);
    DCD_CUDA(cudaMemsetAsync(d_small.p, 0, 64, s));
    launch_gather_flagged(tab.view, n_entries, EF_CODE, perm_a.p, d_small.p, s);
    launches++;
    u32 ncodes = 0;
    u32 n_local_codes = 0;                 // sharded build: codes present on this rank (dense local ids), else unused
    DevBuf<u32> d_glob, d_local_of_glob;   // local id -> global ordinal id and back (0xFFFFFFFF: absent here)
    DCD_CUDA(cudaMemcpyAsync(&ncodes, d_small.p, 4, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    if (ncodes >= (1u << 24)) throw StatusError(DCD_E_CAPACITY, "more than 2^24 distinct codes");
    DevBuf<u16> d_code_chars;
    DevBuf<u32> d_code_str_off(pool, (size_t)ncodes + 1);
    u64 code_chars_total = 0;
    if (ncodes) {
        k_max_len<<<grid_for(ncodes), 256, 0, s>>>(tab.view, perm_a.p, ncodes, d_small.p + 1);
        u32 maxlen = 0;
        DCD_CUDA(cudaMemcpyAsync(&maxlen, d_small.p + 1, 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        launches++;
        DevBuf<u64> ka(pool, ncodes), kb(pool, ncodes);
        DevBuf<u64> d_hist(pool, 8 * 256);
        DevBuf<u8> scratch(pool, sort_scratch_bytes(ncodes));
        const SortPlan full = make_sort_plan(0, 64);
        std::vector<u64> h_hist(8 * 256);
        u32* pa = perm_a.p;
        u32* pb = perm_b.p;
        const int nchunks = (int)((maxlen + 3) / 4);
        for (int ch = nchunks - 1; ch >= 0; ch--) {
            k_chunk_keys<<<grid_for(ncodes), 256, 0, s>>>(tab.view, pa, ncodes, ch, ka.p);
            DCD_CUDA(cudaMemsetAsync(d_hist.p, 0, d_hist.bytes(), s));
            launch_histogram(ka.p, ncodes, full, d_hist.p, s);
            DCD_CUDA(cudaMemcpyAsync(h_hist.data(), d_hist.p, d_hist.bytes(), cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            launches += 2;
            int where = radix_sort_run(ka.p, kb.p, pa, pb, ncodes, full, d_hist.p, h_hist.data(), scratch.p, s, &launches);
            if (where) std::swap(pa, pb);   // keys are rebuilt per chunk, only the permutation carries over
        }
        // ranks -> ent_value ; packed code strings in rank order
        DevBuf<u64> lens(pool, ncodes), lens_incl(pool, ncodes);
        k_assign_ranks<<<grid_for(ncodes), 256, 0, s>>>(tab.view, pa, ncodes, lens.p);
        DevBuf<u64> lb(pool, (ncodes + 2047) / 2048 + 1);
        DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
        launch_inclusive_scan_u64(lens.p, lens_incl.p, ncodes, 0, lb.p, ticket.p, s);
        DCD_CUDA(cudaMemcpyAsync(&code_chars_total, lens_incl.p + (ncodes - 1), 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        d_code_chars = DevBuf<u16>(pool, code_chars_total);
        k_pack_codes<<<grid_for(ncodes), 256, 0, s>>>(tab.view, pa, ncodes, lens_incl.p, d_code_chars.p, d_code_str_off.p);
        launches += 3;
        if (comm && comm->nranks > 1) {
            DevBuf<u64> present(pool, ncodes), present_incl(pool, ncodes);
            k_present<<<grid_for(ncodes), 256, 0, s>>>(tab.view, pa, ncodes, present.p);
            DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
            DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
            launch_inclusive_scan_u64(present.p, present_incl.p, ncodes, 0, lb.p, ticket.p, s);
            u64 nl = 0;
            DCD_CUDA(cudaMemcpyAsync(&nl, present_incl.p + (ncodes - 1), 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            n_local_codes = (u32)nl;
            d_glob = DevBuf<u32>(pool, std::max<u32>(n_local_codes, 1));
            d_local_of_glob = DevBuf<u32>(pool, ncodes);
            k_local_ids<<<grid_for(ncodes), 256, 0, s>>>(tab.view, pa, ncodes, present_incl.p, d_glob.p, d_local_of_glob.p);
            launches += 3;
            DCD_CUDA(cudaStreamSynchronize(s));   // `present*` are released at the end of this scope
        }
        // the key strings travel to the host (what Index.Load builds: SortedList<string,...>, Index.cs:151,341) on the copy
        // stream while expand / sort / encode run
        c->ix.h_code_str_off.resize_uninit((size_t)ncodes + 1);
        c->ix.h_code_chars.resize_uninit(code_chars_total);
        DCD_CUDA(cudaEventRecord(c->copy_event, s));
        DCD_CUDA(cudaStreamWaitEvent(c->copy_stream, c->copy_event, 0));
        DCD_CUDA(cudaMemcpyAsync(c->ix.h_code_str_off.data(), d_code_str_off.p, ((size_t)ncodes + 1) * 4, cudaMemcpyDeviceToHost, c->copy_stream));
        if (code_chars_total)
            DCD_CUDA(cudaMemcpyAsync(c->ix.h_code_chars.data(), d_code_chars.p, code_chars_total * 2, cudaMemcpyDeviceToHost, c->copy_stream));
    }
    launch_codes_to_ids(tab.view, n_words, word_codes.p, s);
    launch_codes_to_ids(tab.view, n_extra, extra_codes.p, s);
    launch_pack_slot_codes(word_slots.p, n_words, word_codes.p, slot_codes.p, s);
    launches += 3;
    DCD_CUDA(cudaGetLastError());
    tm.mark();   // 2: codes done

    // ---- K3: token records -> posting keys ---------------------------------------------------------------
    u64 h64[16];
    const u64 per_tok = (u64)std::min<int>(MAX_CODES, std::max(2, 1 + bp.nvocs));
    const u64 extra_cap = (u64)n_extra * MAX_CODES;
    const u64 key_cap = n_tokens * per_tok + extra_cap + 16;
    DevBuf<u64> keys_a(pool, key_cap), keys_b(pool, key_cap);
    const bool local_ids = d_glob.p != nullptr;
    const int code_bits = bits_for(std::max<u64>(local_ids ? n_local_codes : ncodes, 2));
    const SortPlan plan = make_wide_plan(COORD_BITS, COORD_BITS + code_bits);
    DevBuf<u64> d_hist(pool, 8 * WIDE_HIST_STRIDE);
    DCD_CUDA(cudaMemsetAsync(d_hist.p, 0, d_hist.bytes(), s));
    DCD_CUDA(cudaMemsetAsync(d_err.p, 0, 64, s));
    u64 n_post = 0;
    for (size_t bi = 0; bi < b->batches.size(); bi++) {
        TextBatch& bt = b->batches[bi];
        if (!bt.ntiles) continue;
        DevBuf<u64> lb(pool, (size_t)bt.ntiles + bt.ntiles2 / EX2_GROUP + 2);
        DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
        DCD_CUDA(cudaMemsetAsync(counters64.p, 0, 8, s));
        if (btok[bi].v2) {
            Expand2Params e2{};
            e2.rec = btok[bi].rec.p;
            e2.tile_count = btok[bi].cnt16.p;
            e2.ntiles = bt.ntiles2;
            e2.delta = bt.delta0;
            e2.slot_codes = slot_codes.p;
            e2.overflow = word_codes.p;
            e2.keys = keys_a.p + n_post;
            e2.capacity = key_cap - extra_cap - n_post;
            e2.lookback = lb.p;      // ceil(ntiles2 / 8) groups <= ntiles (8192-unit tiles) + 1
            e2.ticket = ticket.p;
            e2.hist = d_hist.p;
            e2.npasses = plan.npasses;
            for (int i = 0; i < plan.npasses; i++) { e2.pass_shift[i] = plan.shift[i] - COORD_BITS; e2.pass_bits[i] = plan.bits[i]; }
            e2.total = counters64.p;
            e2.error = d_err.p;
            launch_expand2(e2, c->nsm, s);
            launches++;
            u64 got2 = 0;
            DCD_CUDA(cudaMemcpyAsync(&got2, counters64.p, 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            n_post += got2;
            continue;
        }
        ExpandParams ep{};
        ep.tok = btok[bi].tok.p;
        ep.tok_cap = tok_cap;
        ep.tile_count = btok[bi].tile_count.p;
        ep.ntiles = bt.ntiles;
        ep.slot_codes = slot_codes.p;
        ep.overflow = word_codes.p;
        ep.keys = keys_a.p + n_post;
        ep.capacity = key_cap - extra_cap - n_post;
        ep.lookback = lb.p;
        ep.ticket = ticket.p;
        ep.hist = d_hist.p;
        ep.npasses = plan.npasses;
        for (int i = 0; i < plan.npasses; i++) { ep.pass_shift[i] = plan.shift[i] - COORD_BITS; ep.pass_bits[i] = plan.bits[i]; }
        ep.total = counters64.p;
        ep.error = d_err.p;
        launch_expand(ep, c->nsm, s);
        launches++;
        u64 got = 0;
        DCD_CUDA(cudaMemcpyAsync(&got, counters64.p, 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        n_post += got;
    }
    btok.clear();
    u32 herr[16];
    DCD_CUDA(cudaMemcpyAsync(herr, d_err.p, 64, cudaMemcpyDeviceToHost, s));
    DCD_CUDA(cudaStreamSynchronize(s));
    DCD_CUDA(cudaGetLastError());
    if (herr[0]) throw StatusError(DCD_E_CAPACITY, "posting buffer capacity exceeded");
    tm.mark();   // 3: emit done

    // ---- K5: sort by code id (stable; keys are already in coordinate order) ---------------------------
    if (n_post >= (1ull << 32)) throw StatusError(DCD_E_CAPACITY, "more than 2^32 postings in one build: split the corpus into document-range shards");
    DevBuf<u8> sort_scratch(pool, std::max(wide_scratch_bytes(n_post + 1, plan.npasses), sort_scratch_bytes(extra_cap + 1)));
    u64* sorted = keys_a.p;
    u64* other = keys_b.p;
    if (n_post) {
        int where = radix_sort_wide(keys_a.p, keys_b.p, n_post, plan, d_hist.p, sort_scratch.p, s, &launches);
        if (where) std::swap(sorted, other);
    }
    // host-supplied postings: emit, sort on the full key, merge
    u64 n_total = n_post;
    if (n_extra) {
        DevBuf<u64> xa(pool, extra_cap + 1), xb(pool, extra_cap + 1);
        DCD_CUDA(cudaMemsetAsync(d_small.p, 0, 64, s));
        launch_emit_extras(tab.view, extra_codes.p, d_extra_entries.p, d_extra_coord.p, d_extra_kind.p, n_extra, xa.p, d_small.p, s);
        u32 nx = 0;
        DCD_CUDA(cudaMemcpyAsync(&nx, d_small.p, 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        launches++;
        if (nx) {
            const SortPlan full = make_sort_plan(0, 64);
            DevBuf<u64> xh(pool, 8 * 256);
            DCD_CUDA(cudaMemsetAsync(xh.p, 0, xh.bytes(), s));
            launch_histogram(xa.p, nx, full, xh.p, s);
            launches++;
            int where = radix_sort_run(xa.p, xb.p, nullptr, nullptr, nx, full, xh.p, nullptr, sort_scratch.p, s, &launches);
            const u64* xs = where ? xb.p : xa.p;
            launch_merge_small(sorted, n_post, xs, nx, other, s);
            launches++;
            std::swap(sorted, other);
            n_total = n_post + nx;
            DCD_CUDA(cudaStreamSynchronize(s));
        }
    }
    DCD_CUDA(cudaGetLastError());
    tm.mark();   // 4: sort done

    // ---- K6/K7: encode ------------------------------------------------------------------------------------
    ResidentIndex& ix = c->ix;
    ix.ncodes = ncodes;
    ix.coord_origin = b->coord_origin;
    ix.d_code_off = DevBuf<u64>(pool, (size_t)ncodes + 1);
    DCD_CUDA(cudaMemsetAsync(ix.d_code_off.p, 0, ix.d_code_off.bytes(), s));
    u64 arena_total = 0;
    // sharded build: the first delta of each code's segment is taken against the last coordinate of that code on the
    // nearest earlier shard that has it (IndexSequence.Shift / AddRange re-delta of the partial-index stitch,
    // IndexSequence.cs:191-202, Index.cs:660-671) -- all-gather of last_coord[code], SURVEY.md 8e step 4
    DevBuf<u64> first_base;
    xt.mark();   // x2
    if (comm && comm->nranks > 1 && ncodes) {
        DevBuf<u64> my_last(pool, ncodes), all_last(pool, (size_t)ncodes * comm->nranks);
        launch_code_last(sorted, n_total, ncodes, local_ids ? d_local_of_glob.p : nullptr, my_last.p, s);
        comm->allgather(my_last.p, all_last.p, (size_t)ncodes * 8, s);
        first_base = DevBuf<u64>(pool, ncodes);
        launch_first_base(all_last.p, comm->nranks, comm->rank, ncodes, first_base.p, s);
        launches += 2;
        DCD_CUDA(cudaStreamSynchronize(s));   // all_last is released below
    }
    xt.mark();   // x3
    // `other` (the free half of the sort double buffer) receives the ushort stream; the resident index then adopts
    // the front of that block in place (DevPool::shrink) -- no copy
    u16* tmp_arena = reinterpret_cast<u16*>(other);
    const u64 tmp_cap = key_cap * 4;   // 8 bytes per key == 4 ushorts >= the 3 a 40-bit difference can need
    DevBuf<u64> enc_lb;   // look-back words of the encoder / (pipelined) the per-tile prefixes of the lengths pass
    DevBuf<u64> key_base;
    u64 merged_units = 0;
    // Sharded build, peer memory available: the list LENGTHS are computed first (same kernel, nothing stored), the owner split
    // is planned from them, and the stream is then encoded one owner's slice at a time while the previous slice is already
    // being stored into its owner's arena over NVLink (exchange() below).  Every rank takes the same decision: it depends
    // on the communicator and the global dictionary only, never on what this shard holds.
    // Measured at 8 GPUs (profiles/r02_exchange_variants.md): the pipeline hides ~4 ms of the exchange but pays 2.5 ms for the
    // lengths pass and shares the SMs with the scatter -- 34.8 ms against 34.75 ms for encode-then-scatter.  Off by default.
    const char* pipe_env = getenv("DCD_SHARD_PIPELINE");   // read per build; every member of a group must see the same value
    const bool pipelined = comm && comm->nranks > 1 && ncodes && pipe_env && atoi(pipe_env) != 0;
    enc_lb = DevBuf<u64>(pool, pipelined ? encode_lengths_scratch(n_total) : encode_scratch_tiles(n_total) + 1);
    if (n_total) {
        DCD_CUDA(cudaMemsetAsync(enc_lb.p, 0, enc_lb.bytes(), s));
        DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
        DCD_CUDA(cudaMemsetAsync(d_err.p, 0, 64, s));
        if (first_base.p && local_ids && n_local_codes) {
            key_base = DevBuf<u64>(pool, n_local_codes);
            k_base_by_local<<<grid_for(n_local_codes), 256, 0, s>>>(first_base.p, d_glob.p, n_local_codes, key_base.p);
            launches++;
        }
        if (pipelined)
            launch_encode_lengths(sorted, n_total, ncodes, ix.d_code_off.p, first_base.p, local_ids ? d_glob.p : nullptr, enc_lb.p, ticket.p,
                                  counters64.p + 4, counters64.p + 3, s, key_base.p, n_local_codes);
        else
            launch_encode(sorted, n_total, ncodes, tmp_arena, tmp_cap, ix.d_code_off.p, first_base.p, local_ids ? d_glob.p : nullptr, enc_lb.p,
                          ticket.p, counters64.p + 4, counters64.p + 3, d_err.p, s, key_base.p, n_local_codes);
        launches += 1;
        DCD_CUDA(cudaMemcpyAsync(h64, counters64.p, sizeof h64, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaMemcpyAsync(herr, d_err.p, 64, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        DCD_CUDA(cudaGetLastError());
        if (herr[0] || h64[4] > tmp_cap) throw StatusError(DCD_E_CAPACITY, "encoded arena capacity exceeded");
        arena_total = h64[4];
        ix.d_arena.adopt(std::move(other == keys_a.p ? keys_a : keys_b), arena_total);   // same address: tmp_arena stays valid
        ix.max_coord = h64[3];
    }
    ix.arena_units = arena_total;
    // host copy of the list offsets (key table of Index.Load, Index.cs:151,341): on the copy stream, behind the page-table
    // kernels and -- sharded build -- the whole segment exchange
    ix.h_code_off.resize_uninit((size_t)ncodes + 1);
    if (ncodes) {
        DCD_CUDA(cudaEventRecord(c->copy_event, s));
        DCD_CUDA(cudaStreamWaitEvent(c->copy_stream, c->copy_event, 0));
        DCD_CUDA(cudaMemcpyAsync(ix.h_code_off.data(), ix.d_code_off.p, ((size_t)ncodes + 1) * 8, cudaMemcpyDeviceToHost, c->copy_stream));
    }
    const u64* first_base_p = first_base.p;
    ix.d_code_base = std::move(first_base);   // the shard serves queries for its range: the decoder starts every list from its base
    tm.mark();   // 5: encode done

    // ---- pages ----------------------------------------------------------------------------------------------
    {
        u64 npages = 0;
        for (auto& sg : b->segments) { sg.first_ordinal = npages; npages += sg.count; }
        ix.npages = npages;
        ix.d_page_end = DevBuf<u64>(pool, npages);
        size_t ex = 0;
        for (size_t si = 0; si < b->segments.size(); si++) {
            PageSegment& sg = b->segments[si];
            if (!sg.count) continue;
            if (!sg.bulk) {
                DCD_CUDA(cudaMemcpyAsync(ix.d_page_end.p + sg.first_ordinal, b->explicit_page_end.data() + ex, sg.count * 8,
                                         cudaMemcpyHostToDevice, s));
                ex += sg.count;
            }
        }
        for (auto& bt : b->batches) {
            if (bt.segment < 0) continue;
            PageSegment& sg = b->segments[bt.segment];
            if (!sg.count) continue;
            if (sg.identity) {
                DCD_CUDA(cudaMemcpyAsync(ix.d_page_end.p + sg.first_ordinal, bt.coord_end.p, sg.count * 8, cudaMemcpyDeviceToDevice, s));
            } else {
                // drop empty pages (Build.cs:483: they never reach EndPage)
                DevBuf<u64> adv(pool, bt.npages), keep(pool, bt.npages), keep_incl(pool, bt.npages);
                DevBuf<u64> lb(pool, bt.npages / 2048 + 2);
                launch_page_adv(bt.page_off, bt.npages, adv.p, s);
                k_keep_flags<<<grid_for(bt.npages), 256, 0, s>>>(adv.p, bt.npages, keep.p);
                DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
                DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
                launch_inclusive_scan_u64(keep.p, keep_incl.p, bt.npages, 0, lb.p, ticket.p, s);
                sg.d_page_orig = std::make_shared<DevBuf<u32>>(pool, sg.count);
                k_compact_pages<<<grid_for(bt.npages), 256, 0, s>>>(adv.p, keep_incl.p, bt.coord_end.p, bt.npages,
                                                                     ix.d_page_end.p + sg.first_ordinal, sg.d_page_orig->p);
                launches += 4;
                DCD_CUDA(cudaStreamSynchronize(s));
            }
        }
        DCD_CUDA(cudaStreamSynchronize(s));
        ix.segments = std::move(b->segments);
        ix.doc_names = std::move(b->doc_names);
        ix.ndocs = b->ndocs;
    }
    tm.mark();   // 6: pages done

    // ---- sharded build: exchange the encoded segments, owner concatenates --------------
    xt.mark();   // x4
    if (comm) {
        const int R = comm->nranks, me = comm->rank;
        MergedIndex& mg = c->merged;
        mg.rank = me;
        mg.nranks = R;
        ix.global_codes = R > 1;
        // global Index.MaxCoord
        DevBuf<u64> d_mc(pool, 1), d_mc_all(pool, R);
        DCD_CUDA(cudaMemcpyAsync(d_mc.p, &ix.max_coord, 8, cudaMemcpyHostToDevice, s));
        comm->allgather(d_mc.p, d_mc_all.p, 8, s);
        std::vector<u64> mc_all(R);
        DCD_CUDA(cudaMemcpyAsync(mc_all.data(), d_mc_all.p, (size_t)R * 8, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        mg.max_coord = *std::max_element(mc_all.begin(), mc_all.end());
        if (ncodes) {
            const size_t row = (size_t)ncodes + 1;
            DevBuf<u64> all_off(pool, row * R), tot(pool, ncodes), tot_incl(pool, ncodes);
            comm->allgather(ix.d_code_off.p, all_off.p, row * 8, s);
            launch_total_len(all_off.p, R, ncodes, tot.p, s);
            DevBuf<u64> lb(pool, (ncodes + 2047) / 2048 + 1);
            DCD_CUDA(cudaMemsetAsync(lb.p, 0, lb.bytes(), s));
            DCD_CUDA(cudaMemsetAsync(ticket.p, 0, 64, s));
            launch_inclusive_scan_u64(tot.p, tot_incl.p, ncodes, 0, lb.p, ticket.p, s);
            DevBuf<u32> d_splits(pool, R + 1);
            DevBuf<u64> d_cut(pool, (size_t)R * (R + 1));
            launch_find_splits(tot_incl.p, ncodes, R, d_splits.p, s);
            launch_gather_cuts(all_off.p, ncodes, R, d_splits.p, d_cut.p, s);
            launches += 4;
            std::vector<u32> splits(R + 1);
            std::vector<u64> cut((size_t)R * (R + 1));
            std::vector<u64> key_begin(R + 1, 0);   // pipelined: first sorted key of every owner's code range
            DevBuf<u64> d_key_begin(pool, R + 1);
            if (pipelined && n_total) {
                launch_key_begins(sorted, n_total, d_splits.p, R, local_ids ? d_glob.p : nullptr, d_key_begin.p, s);
                launches++;
                DCD_CUDA(cudaMemcpyAsync(key_begin.data(), d_key_begin.p, key_begin.size() * 8, cudaMemcpyDeviceToHost, s));
            }
            // big segments (a few hundred hot lists hold about half of the bytes) go by copy engine: no SM time, full link rate
            const char* ce_env = getenv("DCD_SHARD_CE_MIN");   // ushorts; 0 = everything through the SM scatter (read per build: tests lower it)
            const u64 ce_min = ce_env ? (u64)atoll(ce_env) : 0;   // off: slower than the scatter kernel at 2 and at 8 GPUs (same profile note)
            const u32 big_cap = ce_min ? (u32)(arena_total / ce_min + 16) : 1;
            DevBuf<u64> d_big(pool, 1 + (size_t)big_cap * 4);
            std::vector<u64> big(1 + (size_t)big_cap * 4, 0);
            if (ce_min) {
                launch_big_segments(all_off.p, tot_incl.p, ncodes, R, me, d_splits.p, ce_min, d_big.p, big_cap, s);
                launches++;
                DCD_CUDA(cudaMemcpyAsync(big.data(), d_big.p, big.size() * 8, cudaMemcpyDeviceToHost, s));
            }
            DCD_CUDA(cudaMemcpyAsync(splits.data(), d_splits.p, splits.size() * 4, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaMemcpyAsync(cut.data(), d_cut.p, cut.size() * 8, cudaMemcpyDeviceToHost, s));
            DCD_CUDA(cudaStreamSynchronize(s));
            auto cutv = [&](int r, int o) { return cut[(size_t)r * (R + 1) + o]; };
            mg.code_begin = splits[me];
            mg.code_end = splits[me + 1];
            const u32 n_own = mg.code_end - mg.code_begin;
            for (int r = 0; r < R; r++) merged_units += cutv(r, me + 1) - cutv(r, me);   // my code range on every rank
            mg.units = merged_units;
            mg.d_arena = DevBuf<u16>(pool, merged_units + 8);
            mg.d_off = DevBuf<u64>(pool, (size_t)n_own + 1);
            launch_merged_off(tot_incl.p, mg.code_begin, n_own, mg.d_off.p, s);
            launches++;
            // Fused exchange + merge over NVLink peer memory: every rank stores its segments straight into their final place
            // in the owner's arena.  Fallback (no peer mapping, or DCD_SHARD_P2P=0): NCCL all-to-all, then an owner-side concat.
            std::vector<void*> peers(R, nullptr);
            const char* env = getenv("DCD_SHARD_P2P");
            bool p2p = !(env && env[0] == '0');
            void* slab_base = nullptr;
            for (auto& kv : pool.slabs)
                if ((char*)mg.d_arena.p >= kv.first && (char*)mg.d_arena.p < kv.first + kv.second) slab_base = kv.first;
            p2p = comm->exchange_ptrs(p2p ? (void*)mg.d_arena.p : nullptr, p2p ? slab_base : nullptr, peers.data(), s) && p2p;
            if (p2p) {
                DevBuf<u64> d_peers(pool, R);
                DCD_CUDA(cudaMemcpyAsync(d_peers.p, peers.data(), (size_t)R * 8, cudaMemcpyHostToDevice, s));
                xt.mark();   // x5: plan done
                // copy-engine part: stream k takes the owners k, k + 4, ... behind this rank, so the streams target different owners
                const u64 nbig = ce_min ? std::min<u64>(big[0], big_cap) : 0;
                auto ce_join = [&]() {   // the build stream continues behind the copies
                    for (int k = 0; k < 4 && nbig; k++) {
                        DCD_CUDA(cudaEventRecord(c->xchg_event[60 + k], c->ce_stream[k]));
                        DCD_CUDA(cudaStreamWaitEvent(s, c->xchg_event[60 + k], 0));
                    }
                };
                auto ce_copies = [&]() {
                    if (!nbig) return;
                    DCD_CUDA(cudaEventRecord(c->xchg_event[59], s));   // everything the copies read is encoded
                    for (int k = 0; k < 4; k++) DCD_CUDA(cudaStreamWaitEvent(c->ce_stream[k], c->xchg_event[59], 0));
                    for (int k = 0; k < R; k++) {
                        const int o = (me + 1 + k) % R;
                        for (u64 i = 0; i < nbig; i++) {
                            const u64* ent = big.data() + 1 + i * 4;
                            if ((int)ent[2] != o) continue;
                            DCD_CUDA(cudaMemcpyAsync(reinterpret_cast<u16*>(peers[o]) + ent[3], tmp_arena + ent[0], ent[1] * 2, cudaMemcpyDeviceToDevice,
                                                     c->ce_stream[k & 3]));
                        }
                    }
                };
                DevBuf<u32> tile_codes(pool, scatter_scratch_words(arena_total));
                if (pipelined) {
                    // owner by owner, starting behind this rank (at any time the ranks store to different owners): the slice is
                    // encoded on the build stream and scattered on the exchange stream while the next slice is being encoded
                    launch_scatter_tile_codes(all_off.p + (size_t)me * row, ncodes, arena_total, tile_codes.p, s);
                    launches++;
                    // Step i encodes the i-th part of EVERY owner's key-tile range, then one scatter launch stores those parts to
                    // all owners at once (on the exchange stream, while step i + 1 is being encoded).  A part's ushort range comes
                    // from the per-tile prefixes of the lengths pass.
                    const int PIPE_SUB = 4;
                    struct Piece { u64 ta, tb, u0, u1, c0, c1; };
                    std::vector<Piece> pieces((size_t)PIPE_SUB * R, Piece{0, 0, 0, 0, 0, 0});
                    std::vector<u64> want;   // tile indices whose prefix (units in front of the tile) is needed
                    for (int o = 0; o < R && n_total; o++) {
                        if (key_begin[o + 1] <= key_begin[o]) continue;
                        const u64 ta = key_begin[o] / ENC_TILE, tb = (key_begin[o + 1] + ENC_TILE - 1) / ENC_TILE;
                        for (int i = 0; i < PIPE_SUB; i++) {
                            const u64 a = ta + (tb - ta) * i / PIPE_SUB, e = ta + (tb - ta) * (i + 1) / PIPE_SUB;
                            if (e <= a) continue;
                            Piece pc{a, e, i == 0 ? cutv(me, o) : ~0ull, i == PIPE_SUB - 1 ? cutv(me, o + 1) : ~0ull, cutv(me, o), cutv(me, o + 1)};
                            if (pc.u0 == ~0ull) want.push_back(a);
                            if (pc.u1 == ~0ull) want.push_back(e);
                            pieces[(size_t)i * R + o] = pc;
                        }
                    }
                    std::vector<u64> got(want.size());
                    if (!want.empty()) {
                        DevBuf<u64> d_want(pool, want.size()), d_got(pool, want.size());
                        DCD_CUDA(cudaMemcpyAsync(d_want.p, want.data(), want.size() * 8, cudaMemcpyHostToDevice, s));
                        k_gather_prefix<<<grid_for(want.size()), 256, 0, s>>>(enc_lb.p, d_want.p, want.size(), d_got.p);
                        DCD_CUDA(cudaMemcpyAsync(got.data(), d_got.p, got.size() * 8, cudaMemcpyDeviceToHost, s));
                        DCD_CUDA(cudaStreamSynchronize(s));
                        launches++;
                    }
                    size_t wi = 0;
                    for (int o = 0; o < R; o++)          // (the order `want` was filled in)
                        for (int i = 0; i < PIPE_SUB; i++) {
                            Piece& pc = pieces[(size_t)i * R + o];
                            if (pc.tb <= pc.ta) continue;
                            if (pc.u0 == ~0ull) pc.u0 = got[wi++];
                            if (pc.u1 == ~0ull) pc.u1 = got[wi++];
                            pc.u0 = std::max(pc.u0, pc.c0);   // (an owner with fewer tiles than parts: stay inside its own ushort range)
                            pc.u1 = std::min(pc.u1, pc.c1);
                        }
                    for (int i = 0; i < PIPE_SUB; i++) {
                        u64 r0[16], r1[16];
                        for (int k = 0; k < R; k++) {
                            const int o = (me + 1 + k) % R;
                            const Piece& pc = pieces[(size_t)i * R + o];
                            r0[k] = pc.u0;
                            r1[k] = pc.u1;
                            if (pc.tb <= pc.ta) continue;
                            launch_encode_placed(sorted, n_total, pc.ta, pc.tb, tmp_arena, arena_total, first_base_p, local_ids ? d_glob.p : nullptr,
                                                 enc_lb.p, d_err.p, s, key_base.p);
                            launches++;
                        }
                        DCD_CUDA(cudaEventRecord(c->xchg_event[i], s));
                        DCD_CUDA(cudaStreamWaitEvent(c->xchg_stream, c->xchg_event[i], 0));
                        launch_scatter_ranges(tmp_arena, arena_total, all_off.p, tot_incl.p, ncodes, R, me, d_splits.p,
                                              reinterpret_cast<u16* const*>(d_peers.p), r0, r1, R, 0, tile_codes.p, c->xchg_stream, ce_min);
                        launches++;
                    }
                    ce_copies();
                    ce_join();
                    DCD_CUDA(cudaEventRecord(c->xchg_event[64], c->xchg_stream));
                    DCD_CUDA(cudaStreamWaitEvent(s, c->xchg_event[64], 0));
                } else {
                    u64 r0[16], r1[16];
                    for (int k = 0; k < R; k++) { r0[k] = cutv(me, (me + 1 + k) % R); r1[k] = cutv(me, (me + 1 + k) % R + 1); }
                    ce_copies();   // run beside the scatter kernel: joined behind its launch
                    launch_scatter_tile_codes(all_off.p + (size_t)me * row, ncodes, arena_total, tile_codes.p, s);
                    launch_scatter_ranges(ix.d_arena.p, arena_total, all_off.p, tot_incl.p, ncodes, R, me, d_splits.p,
                                          reinterpret_cast<u16* const*>(d_peers.p), r0, r1, R, 0, tile_codes.p, s, ce_min);
                    ce_join();
                    launches += 2;
                }
                comm->barrier(s);   // every peer has finished storing into this rank's arena
                for (int r = 0; r < R; r++) if (r != me) comm->bytes_moved += (cutv(r, me + 1) - cutv(r, me)) * 2;
                DCD_CUDA(cudaGetLastError());
                c->counters.shard_p2p = 1;
            } else {
                std::vector<size_t> so(R), sn(R), ro(R), rn(R);
                std::vector<u64> slab(R);
                u64 recv_units = 0;
                for (int r = 0; r < R; r++) {
                    so[r] = (size_t)cutv(me, r) * 2;
                    sn[r] = (size_t)(cutv(me, r + 1) - cutv(me, r)) * 2;
                    slab[r] = recv_units;
                    ro[r] = (size_t)recv_units * 2;
                    rn[r] = (size_t)(cutv(r, me + 1) - cutv(r, me)) * 2;
                    recv_units += (cutv(r, me + 1) - cutv(r, me) + 7) & ~7ull;   // 16-byte aligned slabs
                }
                DevBuf<u16> recv(pool, recv_units + 8);
                if (pipelined && n_total) {   // no peer mapping: the whole stream in one go, then the all-to-all
                    launch_encode_placed(sorted, n_total, 0, encode_scratch_tiles(n_total), tmp_arena, arena_total, first_base_p,
                                         local_ids ? d_glob.p : nullptr, enc_lb.p, d_err.p, s, key_base.p);
                    launches++;
                }
                comm->alltoallv(ix.d_arena.p, so.data(), sn.data(), recv.p, ro.data(), rn.data(), s);
                DevBuf<u64> d_slab(pool, R);
                DCD_CUDA(cudaMemcpyAsync(d_slab.p, slab.data(), (size_t)R * 8, cudaMemcpyHostToDevice, s));
                xt.mark();   // x5: plan + all-to-all enqueued
                launch_concat_segments(all_off.p, ncodes, R, mg.code_begin, n_own, mg.d_off.p, recv.p, d_slab.p, mg.d_arena.p, merged_units, s);
                launches++;
                DCD_CUDA(cudaStreamSynchronize(s));
                DCD_CUDA(cudaGetLastError());
                c->counters.shard_p2p = 0;
            }
        }
        mg.valid = true;
    }
    if (xt.n == 5) xt.mark();   // x5 (no codes / no comm)
    xt.mark();   // x6

    // ---- host key table (what Index.Load builds: SortedList<string,...>, Index.cs:151,341) ------------------
    if (!ncodes) {
        ix.h_code_off[0] = 0;
        ix.h_code_str_off.assign(1, 0);
        ix.h_code_chars.resize_uninit(0);
    }
    tm.mark();   // 7
    DCD_CUDA(cudaStreamSynchronize(s));
    DCD_CUDA(cudaStreamSynchronize(c->copy_stream));
    DCD_CUDA(cudaGetLastError());
    ix.valid = true;

    const uint32_t shard_p2p = c->counters.shard_p2p;
    dcd_counters& k = c->counters;
    memset(&k, 0, sizeof k);
    k.text_units = text_units;
    k.tokens = n_tokens;
    k.postings = n_total;
    k.codes = ncodes;
    k.distinct_words = n_words_entries;
    k.encoded_u16 = arena_total;
    k.max_coord = ix.max_coord;
    k.pages = ix.npages;
    k.build_ms_device = tm.ms(0, 7);
    for (int i = 0; i < 7; i++) k.stage_ms[i] = tm.ms(i, i + 1);
    k.kernel_launches = launches;
    k.stage_ms[7] = b->add_ms;
    k.build_ms_wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    k.sort_passes = (uint32_t)plan.npasses;
    k.code_bits = (uint32_t)code_bits;
    if (comm) {
        k.shard_ms[0] = xt.ms(0, 1);
        k.shard_ms[1] = xt.ms(2, 3);
        k.shard_ms[2] = xt.ms(4, 5);   // offsets all-gather, owner split, all-to-all of the encoded segments
        k.shard_ms[3] = xt.ms(5, 6);   // fused peer-memory scatter (or the owner-side concatenation of the fallback)
        k.shard_p2p = shard_p2p;
        k.shard_ranks = (uint32_t)comm->nranks;
        k.shard_rank = (uint32_t)comm->rank;
        k.merged_u16 = merged_units;
        k.exchange_bytes = comm->bytes_moved;
        k.merged_code_begin = c->merged.code_begin;
        k.merged_code_end = c->merged.code_end;
    }
}

__global__ void k_flag_extras(StrTable t, const u32* __restrict__ entries, const u8* __restrict__ kinds, u32 n) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 e = entries[i];
    if (e == ENT_NONE) return;
    atomicOr(t.ent_flags + e, kinds[i] == 1 ? (u32)EF_CODE : (u32)EF_WORD);
}
void launch_flag_extras(StrTable t, const u32* entries, const u8* kinds, u32 n, cudaStream_t s) {
    k_flag_extras<<<grid_for(n, 128), 128, 0, s>>>(t, entries, kinds, n);
}

}  // namespace dcd
