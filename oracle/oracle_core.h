// ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the product.
//
// CPU restatement (plain C++17, no CUDA) of Docodo's index-build and posting-query path, used
// as the parity checker for the CUDA implementation in docodo_b200/csrc and as the timed CPU
// baseline ("port") in bench.py.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load this code.  The product never links it.
//
// The reference (alexarchen/Docodo, C#) cannot be compiled or run in this image (no .NET), so
// this restatement is pinned against the reference's own xUnit known answers instead:
//   XUnitDocodoTest/IndexTest.cs:126-162 (positions), 164-226 (nine counts, wildcard, field),
//   228-270 (#1 count, voc-vs-plain), 272-319 (Builder path); IndexSequenceTest.cs:17-134 (codec);
//   Dict/en.voc as a known-answer set for the English Snowball stemmer (Dict.cs:260-296).
// Russian stemming (StemmersNet 1.1.1 RussianStemmer, not vendored): PARITY UNPINNED.
//
// Every function cites the reference file:line (relative to /root/reference/Docodo.NET) it follows.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

namespace orc {

using u16s = std::u16string;
using u64 = uint64_t;

// ---- character tables (stand-ins for .NET Regex classes / ToLower / ToUpper) ----------------
enum { CLS_L = 1, CLS_N = 2, CLS_W = 4 };
void tables_init();
extern uint8_t g_cls[65536];
extern uint16_t g_lower[65536];
extern uint16_t g_upper[65536];
inline bool is_L(char16_t c) { return g_cls[(uint16_t)c] & CLS_L; }
inline bool is_N(char16_t c) { return g_cls[(uint16_t)c] & CLS_N; }
inline bool is_w(char16_t c) { return g_cls[(uint16_t)c] & CLS_W; }
u16s to_lower(const u16s& s);
u16s to_upper(const u16s& s);
std::string utf16_to_utf8(const u16s& s);   // BinaryWriter.Write(string) payload encoding
u16s utf8_to_utf16(const uint8_t* p, size_t n);

// ---- stemmers (StemmersNet 1.1.1 == Snowball ~2010; restated from the published algorithms) --
u16s stem_english(const u16s& w);   // Porter2; pinned by Dict/en.voc
u16s stem_russian(const u16s& w);   // Snowball Russian; PARITY UNPINNED

// ---- IndexSequence (IndexSequence.cs) --------------------------------------------------------
struct Seq {
    std::vector<uint16_t> self;   // IndexSequence.cs:156
    int R = 0;                    // IndexSequence.cs:163
    std::vector<u64> decode() const;   // SeqEnumerator.MoveNext, IndexSequence.cs:114-136
    size_t count_values() const;       // LINQ Count()
    void shift(u64 s);                 // IndexSequence.cs:191-202
};
struct SeqBuilder {   // IndexSequence.cs:46-102
    Seq seq;
    u64 Last = 0;
    void add(u64 l);
};
Seq op_mul(const Seq& a, const Seq& b);   // IndexSequence.cs:218-284
Seq op_add(const Seq& a, const Seq& b);   // IndexSequence.cs:286-322

// ---- Vocab (Dict.cs:13-104) -------------------------------------------------------------------
struct Vocab {
    std::unordered_map<u16s, int32_t> map;
    std::vector<std::pair<u16s, int32_t>> ordered;   // enumeration order of the Dictionary == insertion (file) order
    char16_t begin = 0, end = 0;   // Range, Dict.cs:92-93
    int stemmer = 0;               // 0 none, 1 en, 2 ru
    bool load(const uint8_t* p, size_t n);   // Dict.cs:71-95
    u16s stem(const u16s& w) const;          // Dict.cs:65
    int search(const u16s& w) const;         // Dict.cs:98-103
};

struct StemmerEntry { int kind; std::vector<std::pair<char16_t, char16_t>> ranges; std::vector<char16_t> singles; };

struct DocPage { int doc = -1; u16s id; u64 coord = 0; };

struct ResultPage { int page_index; u16s id; std::vector<int32_t> pos; float rank() const; };
struct ResultDoc { int doc; u16s name; std::vector<int> pages; float rank = 0; };
struct WordInfo { u16s word; int nFound = 0; };
struct SearchResult {
    bool success = true;
    std::string error;
    std::vector<ResultPage> foundPages;
    std::vector<ResultDoc> foundDocs;   // final order (rank ascending, Search.cs:599)
    std::vector<WordInfo> words;
};

struct Index {
    // configuration
    std::vector<std::shared_ptr<Vocab>> vocs;            // Index.cs:219-224
    std::unordered_set<u16s> stopWords;                  // Index.cs:226-235
    std::vector<StemmerEntry> Stemmers;                  // Index.cs:175-183 (per-index copy)
    int MIN_WORD_LENGTH = 3, MAX_WORD_LENGTH = 32;       // Index.cs:97,113
    int MAX_FOUND_PAGES = 30000, MAX_FOUND_DOCS = 500;   // Index.cs:101-102
    bool bKeepForms = true;                              // Index.cs:154
    Index();

    // ---- builder state (one Builder, canonical MaxDegreeOfParallelism=1; Build.cs:258-437) ----
    std::map<u16s, SeqBuilder> builder;                  // ordinal comparer, Build.cs:266
    u64 maxCoord = 0;                                    // Build.cs:294,326
    u64 coord = 0;                                       // IndexTask local, Build.cs:459
    std::vector<std::pair<u16s, u64>> pages;             // Build.cs:343
    uint64_t totalPostings = 0;

    std::vector<u16s> GetWordCodes(const u16s& word) const;   // Build.cs:175-247 (nullptr case -> flag)
    void AddDoc(const u16s& sourceid, const u16s& name);      // Build.cs:348-356
    void Add(const u16s& code, u64 c);                        // Build.cs:324-341
    void AddWord(const u16s& word, u64 c);                    // Build.cs:302-316
    void EndPage(const u16s& id, u64 c);                      // Build.cs:364-367
    void EndPage(const u16s& id) { EndPage(id, maxCoord); }   // Build.cs:359-362
    void IndexPage(const u16s& id, const u16s& text);         // Build.cs:477-536 (one page of IndexTask)

    // ---- persistence ----
    std::vector<uint8_t> index_bytes, list_bytes;
    void Finish();                                            // Save + MergeIndexes(1 file): Build.cs:370-404, Index.cs:695-752
    bool Load(const uint8_t* ib, size_t in, const uint8_t* lb, size_t ln);   // Index.cs:312-380, Build.cs:119-148

    // ---- loaded index ----
    bool CanSearch = false;
    u64 MaxCoord = 0;
    std::map<u16s, Seq> self;                                 // SortedList ordinal, Index.cs:151
    std::vector<std::pair<u64, DocPage>> PagesList;           // insertion order, Build.cs:41
    std::vector<u16s> DocsList;

    // ---- search ----
    std::vector<u16s> GetLikeWords(const u16s& word) const;   // Search.cs:160-167
    Seq SearchWord(const u16s& word) const;                   // Search.cs:192-260
    Seq SearchField(const u16s& field, const u16s& value) const;   // Search.cs:126-155
    SearchResult Search(const u16s& req, bool has_opt, int dist) const;   // Search.cs:437-617
    SearchResult PrepareSearchResult(const Seq& res, const std::vector<u16s>& filter) const;   // Search.cs:365-420

    // ---- dictionary-side queries over the key table (oracle_aux.cpp) ----
    std::vector<u16s> GetCloseWords(const u16s& word) const;                  // Search.cs:169-174, Levenshtein Index.cs:46-89
    std::vector<u16s> GetSuggessions(const u16s& req, int n) const;           // Search.cs:176-188
    u16s GetWordsGroup(const u16s& hex) const;                                // Index.cs:274-282
    std::vector<std::pair<u16s, int>> CalcHistogram(size_t top) const;        // Index.cs:284-307 (InMemory: Value.Count)
};

// ---- .voc builders (Dict.cs:109-296) ----------------------------------------------------------
struct VocBuilder {
    std::map<u16s, int32_t> dict;          // SortedDictionary, ordinal comparer (Dict.cs:111)
    int stemmer = 0;
    int32_t nG = 1;
    std::map<int32_t, int32_t> replaces;
    void AddWordsGroup(const std::vector<u16s>& grouplist);   // Dict.cs:120-190
    std::vector<uint8_t> build() const;                       // Dict.cs:192-208
};

}  // namespace orc
