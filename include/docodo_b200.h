/* docodo_b200.h -- C ABI of libdocodo_b200.so: B200-native (sm_100a) inverted-index build and
 * posting-list query engine that drops in behind Docodo.NET's `Index` API for that path.
 *
 * Docodo has no plugin/operator ABI; the boundary this library replaces is the public surface
 * of `Docodo.Index` / `Index.Builder` / `IndexSequence` plus the on-disk `.index` /
 * `.index.list` formats.  Each entry point names the reference member it
 * stands in for (file:line relative to /root/reference/Docodo.NET).  The C# shim that binds these
 * with [DllImport] is in csharp/ and INTEGRATION.md.
 *
 * Conventions: every function returns 0 on success or a negative dcd_status; the message is
 * available from dcd_last_error().  Strings are UTF-16LE code units (what a .NET `string` is),
 * never NUL-terminated, always (pointer, length).  Input buffers are owned by the caller and only
 * read during the call unless stated otherwise.  Nothing ever calls back into managed code.
 * There is no CPU fallback: without a CUDA device dcd_ctx_create fails with DCD_E_NO_DEVICE.
 */
#ifndef DOCODO_B200_H
#define DOCODO_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dcd_ctx dcd_ctx;       /* one per Index instance and GPU */
typedef struct dcd_build dcd_build;   /* one Index.Builder / CreateAsync run */

typedef enum {
    DCD_OK = 0,
    DCD_E_INVALID = -1,       /* bad argument */
    DCD_E_NO_DEVICE = -2,     /* no CUDA device / wrong architecture */
    DCD_E_CUDA = -3,          /* CUDA runtime error (message has the detail) */
    DCD_E_CAPACITY = -4,      /* dictionary / posting capacity exceeded (raise dcd_options) */
    DCD_E_FORMAT = -5,        /* malformed .voc / .index / .index.list */
    DCD_E_STATE = -6,         /* call out of order (e.g. search before an index exists) */
    DCD_E_IO = -7,            /* file could not be read / written */
    DCD_E_SYNTAX = -8,        /* "Syntax Error in search request" (Search.cs:503-508) */
    DCD_E_UNSUPPORTED = -9    /* feature outside the implemented path */
} dcd_status;

/* Text encodings of page buffers (IndexPage.text is a UTF-16 .NET string, DataSources.cs:34-42). */
enum { DCD_ENC_LATIN1 = 1 /* one byte == one UTF-16 unit U+0000..U+00FF (ASCII corpora) */,
       DCD_ENC_UTF16 = 2 /* UTF-16LE code units */,
       DCD_ENC_UTF8 = 3 /* well-formed UTF-8 bytes; page offsets are BYTE offsets.  Coordinates stay UTF-16 units
                           (`coord += c.Length`, Build.cs:529-531): a 4-byte sequence counts two (a surrogate pair). */ };

/* Stemmer attached to a vocabulary / the stemmer-only mode (Index.KnownStemmers, Index.cs:175-181). */
enum { DCD_STEM_NONE = 0, DCD_STEM_EN = 1, DCD_STEM_RU = 2 };

/* Tunables: public fields of Docodo.Index (Index.cs:96-115,127,154,183; Build.cs:21). */
typedef struct dcd_options {
    int32_t min_word_length;     /* MIN_WORD_LENGTH = 3  (Index.cs:113) */
    int32_t max_word_length;     /* MAX_WORD_LENGTH = 32 (Index.cs:97); hard limit 48 */
    int32_t max_found_pages;     /* MAX_FOUND_PAGES = 30000 (Index.cs:101) */
    int32_t max_found_docs;      /* MAX_FOUND_DOCS = 500 (Index.cs:102) */
    int32_t keep_forms;          /* bKeepForms = true (Index.cs:154) */
    int32_t stemmers_mask;       /* bit k = KnownStemmers[k] enabled (digit,ru,en,de,fr); 0 == Stemmers.Clear() */
    uint32_t dict_capacity_log2; /* slots of the device string table (default 23 -> 8 Mi slots) */
    uint32_t reserved;
} dcd_options;

typedef struct dcd_counters {   /* Builder.TotalCount / Index.Count / MaxCoord and friends */
    uint64_t text_units;        /* UTF-16 units tokenized */
    uint64_t tokens;            /* tokens with min<=len<=max (before the stop list) */
    uint64_t postings;          /* Builder.Add calls == sum of list lengths */
    uint64_t codes;             /* Index.Count (Index.cs:253) */
    uint64_t distinct_words;
    uint64_t encoded_u16;       /* total ushorts of all IndexSequences */
    uint64_t max_coord;         /* Index.MaxCoord (Index.cs:310) */
    uint64_t pages;             /* entries of IndexPageList */
    double   build_ms_device;   /* device time of the last build (CUDA events) */
    double   stage_ms[8];       /* per stage: tokenize, codes, expand, sort, encode, pages, key table; [7] host add_* */
    uint64_t kernel_launches;   /* kernels launched by the last build / search batch */
    uint32_t sort_passes;       /* radix digit passes of the posting sort in the last build */
    uint32_t code_bits;         /* bits of the code id sorted on */
    double   search_ms_device;  /* device time of the last dcd_search_batch (CUDA events) */
    uint64_t search_hits;       /* coordinates in the final lists of the last batch */
    double   build_ms_wall;     /* host wall time of the last dcd_build_finish; stage_ms[7] = wall time of dcd_build_add_* */
    uint64_t search_bytes;      /* algorithmic bytes of the last batch (2*u16 read + 16*op out + 12*hits) */
    /* sharded build (dcd_build_finish_sharded); zero after a single-builder build */
    double   shard_ms[4];       /* device ms: [0] dictionary unification, [1] first-delta bases, [2] owner split (+ all-to-all in the fallback), [3] fused peer-memory scatter (fallback: owner concat) */
    uint32_t shard_ranks, shard_rank;
    uint64_t merged_u16;        /* ushorts of this rank's code range of the merged index */
    uint64_t exchange_bytes;    /* bytes received from peer ranks during the build */
    uint32_t merged_code_begin, merged_code_end;   /* this rank owns the global ordinal code ids [begin, end) */
    uint32_t shard_p2p;         /* 1: segments were stored straight into the owners' arenas over peer memory; 0: all-to-all + concat */
    uint32_t reserved2;
} dcd_counters;

/* ---- context ------------------------------------------------------------------------------- */
/* new Index(path, InMemory, vocs)  (Index.cs:124-140).  device = CUDA ordinal. */
int32_t dcd_ctx_create(int32_t device, dcd_ctx** out);
void    dcd_ctx_destroy(dcd_ctx* ctx);                 /* Index.Dispose (Index.cs:769-786) */
/* Run every kernel and copy of this context on the caller's CUDA stream (a cudaStream_t; NULL = legacy default
 * stream) instead of the context's private one, so that the caller can order its own work and CUDA events with
 * the library's.  The stream is borrowed, never destroyed.  (No reference counterpart: Docodo has no device.) */
int32_t dcd_ctx_set_stream(dcd_ctx* ctx, void* cuda_stream);
const char* dcd_last_error(dcd_ctx* ctx);              /* ctx may be NULL: last creation error */
int32_t dcd_options_default(dcd_options* o);
int32_t dcd_options_set(dcd_ctx* ctx, const dcd_options* o);
int32_t dcd_counters_get(dcd_ctx* ctx, dcd_counters* out);

/* Index.AddVoc(new Vocab(file))  (Index.cs:221; Dict.cs:31-95): `.voc` bytes = {7-bit-length UTF-8 string, int32}*. */
int32_t dcd_vocab_load(dcd_ctx* ctx, const uint8_t* voc_bytes, size_t n, int32_t stemmer);
/* Hand-made Vocab (IndexTest.cs:242-250): explicit (stem, group) pairs and Range. */
int32_t dcd_vocab_add_pairs(dcd_ctx* ctx, const uint16_t* chars, const uint32_t* offs, const int32_t* groups,
                            int32_t n, uint16_t range_begin, uint16_t range_end, int32_t stemmer);
/* Index.LoadStopWords / AddStopWords (Index.cs:226-235): replaces the set. */
int32_t dcd_stopwords_set(dcd_ctx* ctx, const uint16_t* chars, const uint32_t* offs, int32_t n);

/* ---- build: Index.CreateAsync / IndexTask (Index.cs:424-534, Build.cs:445-564) and Index.Builder (Build.cs:258-437) */
int32_t dcd_build_begin(dcd_ctx* ctx, dcd_build** out);          /* GetBuilder() (Build.cs:443) */
void    dcd_build_abort(dcd_build* b);
/* Document-range sharding across GPUs: this builder indexes a shard whose first character has global
 * coordinate coord_base and whose first bulk document is number doc_number_base -- the coordinate shift that
 * MergeFiles(shift_coords) / IndexSequence.Shift apply when partial indexes are stitched (Index.cs:584-593,671;
 * IndexSequence.cs:191-202), applied up front.  Must precede the first document. */
int32_t dcd_build_set_origin(dcd_build* b, uint64_t coord_base, uint64_t doc_number_base);
/* Builder.AddDoc(sourceid, name) (Build.cs:348-356).  The stored document name is "source:name". */
int32_t dcd_build_add_doc(dcd_build* b, const uint16_t* source, int32_t ns, const uint16_t* name, int32_t nn);
/* One IndexPage of the current document through the page loop of IndexTask (Build.cs:477-536):
 * id "0" is the header page (parsed on the host, Build.cs:485-524), anything else is body text. */
int32_t dcd_build_add_page(dcd_build* b, const uint16_t* id, int32_t nid, const uint16_t* text, size_t n);
/* Bulk form of the same loop for many documents at once.  `text` holds the pages back to back in
 * `encoding`; page p is text[page_offs[p] .. page_offs[p+1]); document d owns pages
 * doc_first_page[d] .. doc_first_page[d+1]-1, whose ids are "1","2",... (no header page);
 * documents are named "<source>:d<ordinal>" with ordinals continuing across calls.
 * DCD_ENC_UTF8: page_offs count bytes; pure-ASCII text is indexed in place as Latin-1, anything else is transcoded on the
 * device to Latin-1 (all code points < U+0100) or UTF-16 first, so the tokenizer and every coordinate see what the reference's
 * decoded .NET string holds (DataSources.cs:357-379).
 * on_device != 0: text, page_offs and doc_first_page are device pointers (text 16-byte aligned) and must stay valid
 * until dcd_build_finish returns; otherwise they are host pointers: page_offs and doc_first_page are copied during
 * the call, the text travels in chunks on a second stream so that dcd_build_finish tokenizes the first chunks while
 * the rest is still on the bus.  Pageable text is staged during the call (the caller may reuse it on return);
 * PAGE-LOCKED text (cudaHostAlloc / cudaHostRegister) is read asynchronously and must stay valid and unmodified
 * until dcd_build_finish (or dcd_build_destroy) returns. */
int32_t dcd_build_add_pages_bulk(dcd_build* b, const void* text, int32_t encoding, const uint64_t* page_offs,
                                 int64_t npages, const uint64_t* doc_first_page, int64_t ndocs,
                                 const uint16_t* source, int32_t ns, int32_t on_device);
/* Builder.AddWord(word, coord) (Build.cs:302-316) and Builder.Add(code, coord) (Build.cs:324-341). */
int32_t dcd_build_add_word(dcd_build* b, const uint16_t* word, int32_t n, uint64_t coord);
int32_t dcd_build_add_posting(dcd_build* b, const uint16_t* code, int32_t n, uint64_t coord);
/* Builder.EndPage(id) / EndPage(id, coord) (Build.cs:359-367). use_coord == 0: coord of the last posting. */
int32_t dcd_build_end_page(dcd_build* b, const uint16_t* id, int32_t nid, int32_t use_coord, uint64_t coord);
/* Builder.Save + MergeIndexes + Load (Build.cs:370-434, Index.cs:488-512): runs the device pipeline and leaves
 * the index resident on the GPU, ready for dcd_search_*.  Consumes `b`. */
int32_t dcd_build_finish(dcd_build* b);

/* ---- document-range sharded build across GPUs ----------------------------------------------------
 * Replaces the partial-index stitch of Index.MergeAll / MergeFiles(shift_coords) / MergeIndexes (Index.cs:537-752) and
 * IndexSequence.Shift (IndexSequence.cs:191-202).  Rank g indexes a contiguous range of documents whose first character
 * has the global coordinate given to dcd_build_set_origin (sum of the text lengths of the ranks before it).  The
 * collective finish unifies the code dictionaries (all-gather), takes each segment's first delta against the previous
 * shard's last coordinate of that code (all-gather), range-partitions the codes over the ranks by merged size and moves
 * the encoded segments to their owner (all-to-all), which concatenates them in rank order.
 * A dcd_comm is one rank's handle of the group.  Two transports: NCCL (one process per GPU; the 128-byte id made by
 * dcd_comm_unique_id on one rank is carried to the others by the host program) and an in-process loopback group for
 * several contexts of one process on one device (each member calls from its own thread). */
typedef struct dcd_comm dcd_comm;
int32_t dcd_comm_unique_id(uint8_t* out128);
int32_t dcd_comm_init_nccl(dcd_ctx* ctx, const uint8_t* id128, int32_t rank, int32_t nranks, dcd_comm** out);
int32_t dcd_comm_init_local(int32_t nranks, dcd_comm** out /* [nranks] */);
void    dcd_comm_destroy(dcd_comm* comm);
/* Collective form of dcd_build_finish: every rank of `comm` calls it.  Afterwards the context's resident index holds this
 * rank's document range with global coordinates and GLOBAL code ids (it answers dcd_search_* for its shard), and the
 * context additionally holds this rank's code range of the merged index.  The shard's lists are kept in the stitched form
 * that travels to the owners (first delta of a list = its first coordinate minus the code's last coordinate on the earlier
 * shards): dcd_search_* add that base back while decoding; dcd_index_sequence / dcd_index_serialize on a shard of rank > 0
 * return the stitched words (use dcd_merged_serialize for Docodo-format files).  Consumes `b`. */
int32_t dcd_build_finish_sharded(dcd_build* b, dcd_comm* comm);
/* This rank's slice of the merged `.index` file: rank 0's slice starts with the u64 MaxCoord header; the concatenation of
 * the slices in rank order is byte-identical to the single-builder file (Build.cs:376-386).  The merged `.index.list` is
 * the concatenation in rank order of the list bytes dcd_index_serialize returns on each rank. */
int32_t dcd_merged_serialized_size(dcd_ctx* ctx, uint64_t* index_bytes);
int32_t dcd_merged_serialize(dcd_ctx* ctx, uint8_t* index_out, uint64_t index_cap);
/* Raw device-to-host copy of this rank's merged code range (e2e benchmark): arena (uint16, counters.merged_u16 entries) and
 * offsets (uint64, merged_code_end - merged_code_begin + 1 entries).  Either pointer may be NULL. */
int32_t dcd_merged_download_raw(dcd_ctx* ctx, uint16_t* arena, uint64_t arena_cap, uint64_t* code_offs, uint64_t offs_cap);

/* ---- persistence: `.index` / `.index.list` (Index.cs:312-380, Build.cs:99-148, 376-386) -------- */
/* Sizes of, then bytes of, the two files for the resident index (exact Docodo byte format). */
int32_t dcd_index_serialized_size(dcd_ctx* ctx, uint64_t* index_bytes, uint64_t* list_bytes);
int32_t dcd_index_serialize(dcd_ctx* ctx, uint8_t* index_out, uint64_t index_cap, uint8_t* list_out, uint64_t list_cap);
int32_t dcd_index_save(dcd_ctx* ctx, const char* dir_utf8);       /* writes <dir>/.index and <dir>/.index.list */
int32_t dcd_index_load_bytes(dcd_ctx* ctx, const uint8_t* index_bytes, uint64_t ni, const uint8_t* list_bytes, uint64_t nl);
int32_t dcd_index_load(dcd_ctx* ctx, const char* dir_utf8);       /* Index.Load (Index.cs:312) */
/* Index.this[key] (Index.cs:255-263): raw ushort stream of one code, valid until the next call on ctx. */
int32_t dcd_index_sequence(dcd_ctx* ctx, const uint16_t* code, int32_t n, const uint16_t** data, int32_t* count);
/* Raw device-to-host copy of the encoded index without file assembly (used by the e2e benchmark):
 * arena (uint16), per-code offsets (uint64, codes+1), page end coordinates (uint64). Any pointer may be NULL. */
int32_t dcd_index_download_raw(dcd_ctx* ctx, uint16_t* arena, uint64_t arena_cap, uint64_t* code_offs, uint64_t* page_ends);

/* ---- query: Index.Search (Search.cs:437-617) ------------------------------------------------------------------- */
/* GetWordCodes (Build.cs:175-247): codes joined by U+0000 into out (capacity in units); returns count via *ncodes. */
int32_t dcd_word_codes(dcd_ctx* ctx, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_units,
                       int32_t* ncodes);

typedef struct dcd_hits dcd_hits;   /* result of a batch; owned by the library until dcd_hits_free */
/* Batch of requests, each evaluated exactly like Index.Search(req, opt): has_opt[i] == 0 -> opt == null (dist 255). */
int32_t dcd_search_batch(dcd_ctx* ctx, const uint16_t* reqs, const uint32_t* req_offs, int32_t nreq,
                         const int32_t* has_opt, const int32_t* dist, dcd_hits** out);
void    dcd_hits_free(dcd_hits* h);
/* Per-request accessors.  status: 0 ok, DCD_E_SYNTAX, DCD_E_STATE ("Index is not built"). */
int32_t dcd_hits_status(const dcd_hits* h, int32_t q);
int32_t dcd_hits_num_pages(const dcd_hits* h, int32_t q);                 /* SearchResult.foundPages.Count */
/* foundPages[k]: ordinal of the page in IndexPageList, rank (ResultDocPage.rank, Search.cs:99-111), positions. */
int32_t dcd_hits_page(const dcd_hits* h, int32_t q, int32_t k, int64_t* page_ordinal, float* rank,
                      const int32_t** pos, int32_t* npos);
/* Flat export of the found pages of every request of the batch (sharded serving: per-shard lists are gathered over NCCL and
 * concatenated in rank order, MAX_FOUND_PAGES applied after the merge).  page_counts[nreq]; ordinals / ranks / npos hold the
 * pages of request 0, 1, ... back to back (capacity `cap` entries; any of them may be NULL); *total = number of pages. */
int32_t dcd_hits_export_pages(const dcd_hits* h, int32_t* page_counts, int64_t* ordinals, float* ranks, int32_t* npos,
                              uint64_t cap, uint64_t* total);
/* NOTE: on a shard of a sharded build dcd_search_batch / dcd_search_batch_pages answer from the shard's lists alone -- exact
 * except for gap-runs of operator* that straddle a shard cut (pages of the documents next to the cut); use the *_sharded
 * forms below for answers identical to the merged index.
 * Sharded serving without host-side result shaping: evaluates the batch like dcd_search_batch but keeps the found pages of
 * every request on the DEVICE as int32 pairs (page ordinal + page_base, positions in page), each request cut after page
 * MAX_FOUND_PAGES + 1 (which keeps one position, Search.cs:409).  page_counts[nreq] and *total (pairs) come back; the pairs
 * are then written request 0, 1, ... back to back by dcd_search_pages_fetch into `pairs` (a device pointer when
 * on_device != 0 -- e.g. the send buffer of the NCCL gather -- else host memory; capacity cap_pairs pairs).  One pending
 * batch per context. */
int32_t dcd_search_batch_pages(dcd_ctx* c, const uint16_t* reqs, const uint32_t* req_offs, int32_t nreq, const int32_t* has_opt,
                               const int32_t* dist, int64_t page_base, int32_t* page_counts, uint64_t* total);
int32_t dcd_search_pages_fetch(dcd_ctx* c, int32_t* pairs, uint64_t cap_pairs, int32_t on_device);
/* EXACT serving from the document-range shards of dcd_build_finish_sharded (query partitioning).  Collective:
 * every rank of `comm` calls it with the same batch.  operator* ignores page and document boundaries (a group ends only on a
 * gap > |R|, IndexSequence.cs:218-284), so a gap-run can straddle a shard cut and the shard-local answers of dcd_search_batch
 * differ from the single-index answer for the documents next to the cut.  Here every rank publishes, per request, the tail of
 * its term lists from the start of the chain (gaps <= the request's largest |R|) that reaches the next shard; that rank puts
 * it in front of its own lists, answers from the chain's first coordinate on, and hands the hits left of its own range back
 * to the rank holding those pages before PrepareSearchResult (Search.cs:365-420) runs.  The per-rank results, concatenated in
 * rank order (MAX_FOUND_PAGES / MAX_FOUND_DOCS applied after the merge), are identical to Index.Search on the merged index:
 * pages, positions and page ranks.  wordInfo.nFound of a rank counts its own coordinate range; sum over ranks. */
int32_t dcd_search_batch_sharded(dcd_ctx* ctx, dcd_comm* comm, const uint16_t* reqs, const uint32_t* req_offs, int32_t nreq,
                                 const int32_t* has_opt, const int32_t* dist, dcd_hits** out);
int32_t dcd_search_batch_pages_sharded(dcd_ctx* c, dcd_comm* comm, const uint16_t* reqs, const uint32_t* req_offs, int32_t nreq,
                                       const int32_t* has_opt, const int32_t* dist, int64_t page_base, int32_t* page_counts,
                                       uint64_t* total);
/* The same export as int32 pairs (page ordinal + page_base, positions in page): the payload of the sharded gather, written
 * in one pass by several host threads.  pairs[2 * cap] (may be page-locked; NULL: only counts and *total). */
int32_t dcd_hits_export_pairs32(const dcd_hits* h, int64_t page_base, int32_t* page_counts, int32_t* pairs, uint64_t cap,
                                uint64_t* total);
/* ResultDocPage.pos of every found page of request q back to back, page order (sum of npos entries; `pos` may be NULL
 * to get only *total).  One call instead of one dcd_hits_page per page when a whole result is compared or shipped. */
int32_t dcd_hits_export_positions(const dcd_hits* h, int32_t q, int32_t* pos, uint64_t cap, uint64_t* total);
int32_t dcd_hits_num_docs(const dcd_hits* h, int32_t q);                  /* SearchResult.foundDocs.Count, rank order */
/* foundDocs[k]: document ordinal, rank (Search.cs:552-557), index range [first,first+n) into foundPages. */
int32_t dcd_hits_doc(const dcd_hits* h, int32_t q, int32_t k, int64_t* doc_ordinal, float* rank, int32_t* first_page,
                     int32_t* npages);
int32_t dcd_hits_num_words(const dcd_hits* h, int32_t q);                 /* SearchResult.words */
int32_t dcd_hits_word(const dcd_hits* h, int32_t q, int32_t k, const uint16_t** word, int32_t* n, int32_t* nfound);
/* Page / document metadata of the resident index (IndexPageList, Build.cs:41-149). */
int32_t dcd_index_page_info(dcd_ctx* ctx, int64_t page_ordinal, const uint16_t** id, int32_t* nid, int64_t* doc_ordinal,
                            uint64_t* start_coord, uint64_t* end_coord);
int32_t dcd_index_doc_name(dcd_ctx* ctx, int64_t doc_ordinal, const uint16_t** name, int32_t* n);

/* ---- dictionary-side queries over the resident key table: one device scan + sort of the scores each.
 * Results are strings joined by U+0000 into `out` (capacity `cap` units); *out_units / *count receive the total length and
 * the number of strings.  out == NULL: sizes only. */
/* GetLikeWords (Search.cs:160-167): '_' stands for any run of characters; keys starting with a letter, first 100 in key order. */
int32_t dcd_like_words(dcd_ctx* ctx, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_units, int32_t* count);
/* GetCloseWords (Search.cs:169-174): the 10 keys nearest to `word` by Levenshtein distance (Index.cs:46-89), ties in key order. */
int32_t dcd_close_words(dcd_ctx* ctx, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_units, int32_t* count);
/* GetSuggessions (Search.cs:176-188): completions of the request's last word, most frequent first (by sequence length). */
int32_t dcd_suggestions(dcd_ctx* ctx, const uint16_t* req, int32_t n, int32_t max_words, uint16_t* out, int32_t cap,
                        int32_t* out_units, int32_t* count);
/* CalcHistogram (Index.cs:284-307) of an in-memory index: the `top` (reference: 1000) keys with the longest sequences and
 * their lengths in ushorts; "#..." keys are shown as "(" + the first 20 words of the vocabulary group + ")". */
int32_t dcd_histogram(dcd_ctx* ctx, int32_t top, uint16_t* out, int32_t cap, int32_t* out_units, uint64_t* counts, int32_t* count);

/* ---- IndexSequence algebra on caller-supplied lists (IndexSequence.cs:63-84,114-136,191-322) ------------------- */
/* Encode ascending values on the device (Builder.Add): out_cap in ushorts; *out_n receives the count. */
int32_t dcd_seq_encode(dcd_ctx* ctx, const uint64_t* values, uint64_t n, uint16_t* out, uint64_t out_cap, uint64_t* out_n);
int32_t dcd_seq_decode(dcd_ctx* ctx, const uint16_t* words, uint64_t n, uint64_t* out, uint64_t out_cap, uint64_t* out_n);
/* op '*' (0) or '+' (1) of two encoded lists with their R; result encoded + R. */
int32_t dcd_seq_op(dcd_ctx* ctx, int32_t op, const uint16_t* a, uint64_t na, int32_t Ra, const uint16_t* b, uint64_t nb,
                   int32_t Rb, uint16_t* out, uint64_t out_cap, uint64_t* out_n, int32_t* Rout);

/* ---- allocation-free entry point ------------------------------------------------------------------------------ */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same stream as dcd_seq_encode (ascending
 * values -> Builder.Add's ushort difference stream), but every pointer is a DEVICE pointer, the work is only enqueued on the
 * context's stream (no synchronisation, no host copy), and nothing is allocated or freed: the encoder's look-back scratch
 * is a fixed 64 KiB __device__ array inside the library, which limits n to DCD_NOALLOC_MAX_VALUES.  Values may use all 64 bits.
 *   d_out_n[0] = ushorts written; d_out_n[1] = 0 ok, 1 out_cap too small (stream incomplete), 2 values not ascending.
 * n > DCD_NOALLOC_MAX_VALUES returns DCD_E_CAPACITY (dcd_last_error has the text) and enqueues nothing.
 * All calls on one device share the scratch array: calls on the same stream are safe back to back (each clears what it
 * reads), calls that overlap on the GPU (different streams without a dependency) race on it. */
#define DCD_NOALLOC_MAX_VALUES (1ull << 24)
int32_t dcd_seq_encode_noalloc(dcd_ctx* ctx, const uint64_t* d_values, uint64_t n, uint16_t* d_out, uint64_t out_cap,
                               uint64_t* d_out_n /* [2] */);

/* ---- synthetic corpus (BASELINE.json config 3/4 generator; benchmark utility, not a Docodo API) ------------- */
/* Generates ndocs documents x pages_per_doc pages of Zipf(s) text over `words` (packed UTF-16/ASCII vocabulary) on
 * the device.  Returns device pointers owned by ctx (freed by dcd_synth_free / ctx destroy). */
typedef struct dcd_synth {
    void* d_text;               /* DCD_ENC_LATIN1 bytes */
    uint64_t* d_page_offs;      /* npages + 1 */
    uint64_t* d_doc_first_page; /* ndocs + 1 */
    uint64_t text_bytes, npages, ndocs;
} dcd_synth;
int32_t dcd_synth_generate(dcd_ctx* ctx, const uint8_t* words, const uint32_t* word_offs, int32_t nwords, double zipf_s,
                           uint64_t seed, int64_t ndocs, int32_t pages_per_doc, int32_t min_page_chars, int64_t first_doc,
                           dcd_synth* out);
int32_t dcd_synth_free(dcd_ctx* ctx, dcd_synth* s);
int32_t dcd_memcpy_d2h(dcd_ctx* ctx, void* dst, const void* d_src, uint64_t n);
/* Host evaluation of the stemmer source shared with the device kernel (IStemmer.Stem, Dict.cs:65); needs no GPU. */
int32_t dcd_stem_host(int32_t stemmer, const uint16_t* word, int32_t n, uint16_t* out, int32_t cap, int32_t* out_n);

#ifdef __cplusplus
}
#endif
#endif /* DOCODO_B200_H */
