// K8/K9: batched evaluation of posting-list expressions on the device.
//
// Every query is a small postfix program over a stack of coordinate lists:
//   LOAD   n codes   push  OR (operator+, IndexSequence.cs:286-322) of the decoded lists of n codes
//   MUL    absR,ord  pop b, pop a, push a*b   (proximity AND with grouping, IndexSequence.cs:218-284)
//   ADD              pop b, pop a, push a+b
//   COUNT  slot      nfound[slot] = |top|     (SearchSequence.d(): wordInfo.nFound, Search.cs:306-307)
// R values are data independent, so the host resolves absR / ordered per operator while planning.
// One CTA interprets one program; all lists live in a per-query scratch region of HBM (L2 resident for typical sizes).
// The final list is mapped to (page ordinal, position) by an upper_bound over the page end coordinates
// (IndexPageList.GetPage, Build.cs:74-97; PrepareSearchResult, Search.cs:365-420), grouped by page and ranked
// (ResultDocPage.rank, Search.cs:99-111).
#pragma once
#include "common.cuh"

namespace dcd {

enum : u32 { QOP_LOAD = 0, QOP_MUL = 1, QOP_ADD = 2, QOP_COUNT = 3, QOP_LOADC = 4 /* LOAD of one code whose decoded list is cached */ };

struct QInstr {
    u32 op;
    u32 a;         // LOAD: first index into the code-id array; MUL: absR; COUNT: nfound slot
    u32 b;         // LOAD: number of codes; MUL: ordered (0/1)
    u32 cap;       // LOAD with b > 1: the output region is two halves of `cap` elements (ping-pong of the OR fold)
    u64 out_off;   // LOAD/MUL/ADD: output offset (u64 elements) inside the query's list area
    u64 src;       // LOADC: device address of the decoded list (cap = its length); the list area is only used under a halo
    u32 lim_cap;   // LOADC in a limited program: elements the planned operator outputs have room for
    u32 pad;
};

struct QDesc {
    u32 instr_begin, instr_end;
    u64 list_off;    // start of this query's list area (u64 elements) in the global list scratch
    u64 tmp_off;     // start of this query's temp area; tmp_cap elements per temp array
    u64 tmp_cap;
    u64 hit_off;     // start of this query's region in the hit arrays
    u64 hit_cap;
    // Early termination (the reference shapes at most MAX_FOUND_PAGES + 1 pages, Search.cs:409): every LOAD decodes its list
    // only up to coordinate `limit` (at most in.cap elements); the final list is trusted below the start of the last chain
    // of the union of the leaf lists (gaps <= rmax) in front of `limit` -- a run of any operator* that reaches `limit` may
    // go on beyond it.  When that does not yield max_pages + 1 pages the program is flagged in QBatch::incomplete and the
    // host runs it again with a larger limit.  ~0: no limit (plain evaluation).
    u64 limit;
    u64 rmax;
};

struct QBatch {
    const QDesc* desc;
    const QInstr* instr;
    const u32* code_ids;
    u32 nqueries;
    // index
    const u16* arena;
    const u64* code_off;
    const u64* code_base;   // nullable; shard of a sharded build: the lists' first deltas are taken against this per-code coordinate
    const u64* page_end;
    u64 npages;
    u64 coord_origin;  // start coordinate of the first page (0 unless the index is a document-range shard)
    u32 max_pages;     // MAX_FOUND_PAGES: results are cut after page max_pages + 1 (PrepareSearchResult stops there, Search.cs:409)
    // scratch
    u64* lists;        // list areas
    u64* tmp_dec;      // decode temp     [sum tmp_cap]
    u8* tmp_tag;       // suffix flags of operator* (1 byte per union element) [sum tmp_cap]
    // outputs
    u32* nfound;       // per term slot
    u32* hit_page;     // [sum hit_cap]
    int* hit_pos;
    u32* page_first;   // first hit index (relative to the query's region) of each found page
    float* page_rank;
    u32* nhits;        // per query
    u32* npages_found; // per query
    u32* error;        // [0] stack/capacity violation
    u32 gallop_ratio;  // operator*: galloping when the longer list has at least this many times the elements of the shorter (0: never)
    u32* incomplete;   // nullable; [nqueries] 1: a limited program did not reach max_pages + 1 pages (or a LOAD ran out of room)
    unsigned long long* cta_ns;   // nullable: per-program execution time in ns (development trace)
    // ---- exact serving from document-range shards (null / 0 otherwise) ---------------------------------------
    // operator* ignores document boundaries, so a gap-run can straddle a shard cut.  Every run is a sub-chain of the union U
    // of the program's leaf lists with gaps <= rmax; the chain of U that bridges this shard's LEFT cut is completed by a HALO:
    // the elements of that chain held by the ranks to the left, prepended to the LOAD results.  The shard's answer is then
    // exact on [first halo element, keep_below), where keep_below = start of the chain that bridges its RIGHT cut (the next
    // shard answers from there on).
    const u32* load_base;     // [nqueries] first LOAD slot of the program; its k-th LOAD uses slot load_base + k
    const u32* halo_cnt;      // [load slots] halo elements of the LOAD (0: none)
    const u64* halo_off;      // [load slots] start in halo_vals
    const u64* halo_vals;
    const u64* keep_below;    // [nqueries] final elements >= this are dropped (~0: keep all)
    u32 defer_pages;          // 1: no hits -> pages step; fin_* receive the final list (k_query_pages runs it after the exchange)
    u64** fin_ptr;            // [nqueries] final list (inside the program's list area)
    u32* fin_n;               // [nqueries] elements kept (< keep_below)
    u32* fin_foreign;         // [nqueries] how many of them lie left of coord_origin (they belong to earlier shards)
};

// Halo pass of exact sharded serving: runs only the LOADs of every program and finds the chain of U that ends at the
// shard's last element.  Per program: HaloDesc; per LOAD slot: the tail of the list from the chain start on (only when the
// chain reaches the next shard: next_first - last <= rmax), appended to `vals` through the atomic cursor.
struct HaloDesc {
    u64 lo;        // first element of the chain that ends at the last element of U on this shard
    u64 first;     // first element of U on this shard
    u64 last;
    u32 flags;     // bit 0: U non-empty, 1: bridged to a later shard, 2: the chain covers all of U on this shard
    u32 pad;
};
struct HaloPass {
    const u64* rmax;          // [nqueries] largest |R| of the program (chain gap bound)
    const u64* next_first;    // [ncodes] first coordinate of the code on the LATER shards (~0: none)
    HaloDesc* desc;           // [nqueries]
    u32* slot_cnt;            // [load slots]
    u64* slot_off;            // [load slots] start of the slot's tail in vals
    u64* vals;
    u64 vals_cap;
    unsigned long long* cursor;   // elements used in vals
};
void launch_query_halo(const QBatch& qb, const HaloPass& hp, cudaStream_t s);

// Hits -> pages of exact sharded serving, after the exchange: program q's final list = its own kept elements followed by
// the elements the later ranks found inside this shard's coordinate range [coord_origin, range_end).
struct PagesPass {
    const u64* const* own_ptr;   // own kept elements of program q, own_n[q] of them
    const u32* own_n;
    int nranks, rank;
    const u64* const* f_vals; // [nranks] device pointers: rank r's foreign elements (all programs)
    const u64* const* f_off;  // [nranks] -> [nqueries] start of program q
    const u32* const* f_cnt;  // [nranks] -> [nqueries]
    u64 range_end;            // coordinate where the next shard starts (~0 for the last rank)
    u64* work;                // assembled lists, program q at desc[q].list_off (capacity desc[q].hit_cap)
};
void launch_query_pages(const QBatch& qb, const PagesPass& pp, cudaStream_t s);
// own part / foreign part of the deferred final lists -> dense arrays
void launch_pack_final(const QBatch& qb, const u64* own_off, const u64* foreign_off, u64* own_vals, u64* foreign_vals, cudaStream_t s);

// Decoded-list cache: the lists of njobs codes decoded by all SMs at once (4096-word chunks: partial sums, per-list scan of the
// chunk totals, decode with the chunk's base).  job arrays on the device; chunk c belongs to job chunk_job[c], is its
// chunk_idx[c]-th chunk.  psum / pcnt: nchunks scratch each.
struct DecodeJobs {
    const u64* word_off;    // [njobs] start of the list in the arena
    const u64* nwords;      // [njobs]
    const u64* base;        // [njobs] value the running coordinate starts from (first-delta base of a shard, else 0)
    u64* const* out;        // [njobs] destination (nvals elements)
    const u32* chunk_job;   // [nchunks]
    const u32* chunk_idx;   // [nchunks]
    const u32* job_first_chunk;   // [njobs + 1]
    u32 njobs, nchunks;
    u64* psum;
    u32* pcnt;
};
void launch_decode_jobs(const u16* arena, const DecodeJobs& dj, cudaStream_t s);
// nvals[c] = number of coordinates in the list of code c (terminator words of its ushort stream): wordInfo.nFound of a
// single-code term without decoding the list
void launch_code_nvals(const u16* arena, const u64* code_off, u32 ncodes, u32* nvals, cudaStream_t s);
// dst[seg_dst[i] .. + seg_cnt[i]) = src[seg_src[i] ..) for nseg segments (src may be NULL: seg_src are then absolute element addresses)
void launch_copy_segments(const u64* src, u64* dst, const u64* seg_src, const u64* seg_dst, const u32* seg_cnt, u32 nseg, cudaStream_t s);
// first[c] = first coordinate of code c in this index (base[c] + the list's first value), ~0 when the list is empty
void launch_first_coords(const u16* arena, const u64* code_off, const u64* code_base, u32 ncodes, u64* first, cudaStream_t s);
// all[r][0 .. row): out[c] = min over ranks r > me of all[r][c] for c < ncodes (~0 = none)
void launch_next_first(const u64* all, int nranks, int me, u32 row, u32 ncodes, u64* out, cudaStream_t s);

void launch_query_batch(const QBatch& qb, cudaStream_t s);

// dense packing of per-query results: offsets = exclusive prefix of counts (computed by the caller)
void launch_pack_results(const QBatch& qb, const u64* hit_dense_off, const u64* page_dense_off, u32* d_hit_page, int* d_hit_pos,
                         u32* d_page_first, float* d_page_rank, cudaStream_t s);

// found pages of every program as int32 pairs (page ordinal + page_base, positions in page; the page after MAX_FOUND_PAGES keeps
// one position, Search.cs:409) at pair offset page_dense_off[program]
void launch_pack_pairs(const QBatch& qb, const u64* page_dense_off, long long page_base, int* out_pairs, cudaStream_t s);
// request r: cnt[r] pairs from device address src[r] to out + 2 * dst_off[r]
void launch_gather_pairs(const u64* src, const u64* dst_off, const u32* cnt, u32 nreq, int* out, cudaStream_t s);

// Stand-alone list algebra for dcd_seq_* (single CTA): see capi.
void launch_seq_op(int op, const u16* a, u64 na, const u16* b, u64 nb, u32 absR, u32 ordered, u64* va, u64* vb, u8* tmp_tag,
                   u64* scratch /* na + nb u64 */, u32 gallop_ratio, u64* out, u32* out_n, cudaStream_t s);
void launch_seq_decode(const u16* w, u64 n, u64* out, u32* out_n, cudaStream_t s);

}  // namespace dcd
