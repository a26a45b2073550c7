// ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle_core.h for the header comment and pinning).
#include "oracle_core.h"

#include <algorithm>
#include <cmath>
#include <cstring>

#include "unicode_tables.h"

namespace orc {

// =====================================================================================
// character tables
// =====================================================================================
uint8_t g_cls[65536];
uint16_t g_lower[65536];
uint16_t g_upper[65536];
static bool g_tables_ready = false;

void tables_init() {
    if (g_tables_ready) return;
    memset(g_cls, 0, sizeof g_cls);
    for (int i = 0; i < ORC_N_CLS_RUNS; i++)
        for (uint32_t c = ORC_CLS_RUNS[i][0]; c <= ORC_CLS_RUNS[i][1]; c++) g_cls[c] = (uint8_t)ORC_CLS_RUNS[i][2];
    for (uint32_t c = 0; c < 65536; c++) g_lower[c] = g_upper[c] = (uint16_t)c;
    for (int i = 0; i < ORC_N_LOWER_PAIRS; i++) g_lower[ORC_LOWER_PAIRS[i][0]] = ORC_LOWER_PAIRS[i][1];
    for (int i = 0; i < ORC_N_UPPER_PAIRS; i++) g_upper[ORC_UPPER_PAIRS[i][0]] = ORC_UPPER_PAIRS[i][1];
    g_tables_ready = true;
}

u16s to_lower(const u16s& s) {
    u16s r = s;
    for (auto& c : r) c = (char16_t)g_lower[(uint16_t)c];
    return r;
}
u16s to_upper(const u16s& s) {
    u16s r = s;
    for (auto& c : r) c = (char16_t)g_upper[(uint16_t)c];
    return r;
}

// .NET UTF8Encoding (replacement fallback): lone surrogates become U+FFFD.
std::string utf16_to_utf8(const u16s& s) {
    std::string o;
    for (size_t i = 0; i < s.size(); i++) {
        uint32_t c = (uint16_t)s[i];
        if (c >= 0xD800 && c <= 0xDBFF && i + 1 < s.size() && (uint16_t)s[i + 1] >= 0xDC00 && (uint16_t)s[i + 1] <= 0xDFFF) {
            c = 0x10000 + ((c - 0xD800) << 10) + ((uint16_t)s[i + 1] - 0xDC00);
            i++;
        } else if (c >= 0xD800 && c <= 0xDFFF) {
            c = 0xFFFD;
        }
        if (c < 0x80) o.push_back((char)c);
        else if (c < 0x800) { o.push_back((char)(0xC0 | (c >> 6))); o.push_back((char)(0x80 | (c & 0x3F))); }
        else if (c < 0x10000) {
            o.push_back((char)(0xE0 | (c >> 12))); o.push_back((char)(0x80 | ((c >> 6) & 0x3F))); o.push_back((char)(0x80 | (c & 0x3F)));
        } else {
            o.push_back((char)(0xF0 | (c >> 18))); o.push_back((char)(0x80 | ((c >> 12) & 0x3F)));
            o.push_back((char)(0x80 | ((c >> 6) & 0x3F))); o.push_back((char)(0x80 | (c & 0x3F)));
        }
    }
    return o;
}

u16s utf8_to_utf16(const uint8_t* p, size_t n) {
    u16s o;
    size_t i = 0;
    while (i < n) {
        uint32_t c = p[i];
        int extra = 0;
        if (c < 0x80) extra = 0;
        else if ((c & 0xE0) == 0xC0) { c &= 0x1F; extra = 1; }
        else if ((c & 0xF0) == 0xE0) { c &= 0x0F; extra = 2; }
        else if ((c & 0xF8) == 0xF0) { c &= 0x07; extra = 3; }
        else { o.push_back(0xFFFD); i++; continue; }
        if (extra > 0 && i + extra >= n) { o.push_back(0xFFFD); i++; continue; }
        bool ok = true;
        for (int k = 1; k <= extra; k++) {
            if ((p[i + k] & 0xC0) != 0x80) { ok = false; break; }
            c = (c << 6) | (p[i + k] & 0x3F);
        }
        if (!ok) { o.push_back(0xFFFD); i++; continue; }
        i += extra + 1;
        if (c >= 0x10000) {
            c -= 0x10000;
            o.push_back((char16_t)(0xD800 + (c >> 10)));
            o.push_back((char16_t)(0xDC00 + (c & 0x3FF)));
        } else {
            o.push_back((char16_t)c);
        }
    }
    return o;
}

// =====================================================================================
// IndexSequence  (IndexSequence.cs)
// =====================================================================================
static const int BITS = 15;                 // IndexSequence.cs:13
static const uint16_t OVERFLOW_ = 1 << 15;   // IndexSequence.cs:14
static const uint16_t MASK = OVERFLOW_ - 1;  // IndexSequence.cs:15

// IndexSequence.Builder.Add, IndexSequence.cs:63-84 (min == 0 on the path)
void SeqBuilder::add(u64 l) {
    u64 diff = l - Last;
    do {
        if (diff > MASK) {
            seq.self.push_back((uint16_t)(OVERFLOW_ | (diff & MASK)));
            diff >>= BITS;
        } else {
            seq.self.push_back((uint16_t)diff);
            break;
        }
    } while (diff > 0);
    Last = l;
}

// SeqEnumerator, IndexSequence.cs:104-154
struct SeqEnum {
    const Seq* seq;
    size_t i = 0;
    u64 Last = 0;
    explicit SeqEnum(const Seq& s) : seq(&s) {}
    bool MoveNext() {
        bool bNeedMore = false;
        int shift = 0;
        do {
            if (i < seq->self.size()) {
                uint16_t value = seq->self[i++];
                bNeedMore = (value & OVERFLOW_) != 0;
                Last += shift < 64 ? (((u64)(value & MASK)) << shift) : 0;
                shift += BITS;
            } else {
                return false;
            }
        } while (bNeedMore);
        return true;
    }
    u64 Current() const { return Last; }
};

std::vector<u64> Seq::decode() const {
    std::vector<u64> v;
    SeqEnum e(*this);
    while (e.MoveNext()) v.push_back(e.Current());
    return v;
}
size_t Seq::count_values() const {
    size_t n = 0;
    SeqEnum e(*this);
    while (e.MoveNext()) n++;
    return n;
}

// IndexSequence.Shift, IndexSequence.cs:191-202
void Seq::shift(u64 s) {
    if (self.empty() || s == 0) return;
    SeqEnum en(*this);
    en.MoveNext();
    SeqBuilder b;
    b.add(s + en.Current());
    size_t q = 0;
    do {
    } while ((self[q++] & OVERFLOW_) != 0);
    self.erase(self.begin(), self.begin() + q);
    self.insert(self.begin(), b.seq.self.begin(), b.seq.self.end());
}

// operator *, IndexSequence.cs:218-284 -- restated statement by statement.
Seq op_mul(const Seq& seq1, const Seq& seq2) {
    int absR = std::max(std::abs(seq1.R), std::abs(seq2.R));
    int R = ((seq1.R < 0) && (seq2.R < 0)) ? -absR : absR;
    SeqBuilder newSeq;
    newSeq.seq.R = R;
    std::vector<u64> newGroup;
    SeqEnum enums[2] = {SeqEnum(seq1), SeqEnum(seq2)};
    bool move[2] = {true, true};
    bool can[2] = {true, true};
    bool IsInGr[2] = {false, false};
    do {
        u64 valToAdd = 0;
        if (!move[0] && !move[1]) break;
        if (move[0] && can[0]) can[0] = enums[0].MoveNext();
        if (move[1] && can[1]) can[1] = enums[1].MoveNext();
        move[0] = move[1] = false;
        if (!can[0] && !can[1]) break;
        if (!can[0]) { valToAdd = enums[1].Current(); move[1] = true; }
        else if (!can[1]) { valToAdd = enums[0].Current(); move[0] = true; }
        else if (enums[0].Current() < enums[1].Current()) { valToAdd = enums[0].Current(); move[0] = true; }
        else if (enums[0].Current() > enums[1].Current()) { valToAdd = enums[1].Current(); move[1] = true; }
        else { move[0] = move[1] = true; valToAdd = enums[1].Current(); }

        if (!newGroup.empty()) {
            bool bfinish = false;
            if ((absR != 0) && (valToAdd - newGroup.back() > (u64)absR)) bfinish = true;
            else if ((R < 0) && move[0] && !IsInGr[0]) bfinish = true;
            if (bfinish) {
                if (IsInGr[0] && IsInGr[1])
                    for (u64 v : newGroup) newSeq.add(v);
                IsInGr[0] = IsInGr[1] = false;
                newGroup.clear();
                if (!can[0] || !can[1]) break;
            }
        }
        if (move[0]) IsInGr[0] = true;
        if (move[1]) IsInGr[1] = true;
        newGroup.push_back(valToAdd);
    } while (can[0] || can[1]);
    if (IsInGr[0] && IsInGr[1])
        for (u64 v : newGroup) newSeq.add(v);
    return newSeq.seq;
}

// operator +, IndexSequence.cs:286-322
Seq op_add(const Seq& seq1, const Seq& seq2) {
    int absR = std::max(std::abs(seq1.R), std::abs(seq2.R));
    int R = ((seq1.R < 0) && (seq2.R < 0)) ? -absR : absR;
    SeqBuilder res;
    res.seq.R = R;
    SeqEnum seqe[2] = {SeqEnum(seq1), SeqEnum(seq2)};
    bool Move[2] = {true, true};
    bool Exists[2] = {false, false};
    do {
        for (int q = 0; q < 2; q++) {
            if (Move[q]) Exists[q] = seqe[q].MoveNext();
            Move[q] = false;
        }
        if (Exists[0] || Exists[1]) {
            if (!Exists[1]) { res.add(seqe[0].Current()); Move[0] = true; }
            else if (!Exists[0]) { res.add(seqe[1].Current()); Move[1] = true; }
            else {
                if (seqe[0].Current() <= seqe[1].Current()) Move[0] = true;
                if (seqe[1].Current() <= seqe[0].Current()) Move[1] = true;
                if (Move[0]) res.add(seqe[0].Current());
                else res.add(seqe[1].Current());
            }
        }
    } while (Exists[0] || Exists[1]);
    return res.seq;
}

// =====================================================================================
// Vocab  (Dict.cs:13-104)
// =====================================================================================
// BinaryReader.ReadString = 7-bit encoded length + UTF-8 bytes; then ReadInt32 (Dict.cs:79).
bool Vocab::load(const uint8_t* p, size_t n) {
    map.clear();
    ordered.clear();
    size_t i = 0;
    u16s first_ge_a, last;
    bool have_first = false;
    while (i < n) {
        uint32_t len = 0;
        int sh = 0;
        while (i < n) {
            uint8_t b = p[i++];
            len |= (uint32_t)(b & 0x7F) << sh;
            sh += 7;
            if (!(b & 0x80)) break;
        }
        if (i + len + 4 > n) break;   // EndOfStreamException ends the loop (Dict.cs:84)
        u16s key = utf8_to_utf16(p + i, len);
        i += len;
        int32_t v;
        memcpy(&v, p + i, 4);
        i += 4;
        if (map.count(key)) return false;   // Dictionary.Add would throw
        map.emplace(key, v);
        ordered.emplace_back(key, v);
        if (!have_first && !key.empty() && key[0] >= u'a') { first_ge_a = key; have_first = true; }
        last = key;
    }
    // Dict.cs:92-93 -- enumeration order of a never-shrunk Dictionary is insertion (file) order.
    if (!have_first || last.empty()) return false;
    begin = first_ge_a[0];
    end = last[0];
    return true;
}
u16s Vocab::stem(const u16s& w) const {
    if (stemmer == 1) return stem_english(w);
    if (stemmer == 2) return stem_russian(w);
    return w;
}
int Vocab::search(const u16s& w) const {
    auto it = map.find(w);
    return it == map.end() ? 0 : it->second;
}

// =====================================================================================
// Index: configuration
// =====================================================================================
// Index.KnownStemmers, Index.cs:175-181.  kind: 0 = null stemmer ("digit"), 1 en, 2 ru,
// 3 de, 4 fr (German/French Snowball are NOT restated: words whose characters all fall in
// those ranges but not in a-z keep `stemmed == firststemmed`; documented gap in DESIGN.md).
Index::Index() {
    tables_init();
    Stemmers.push_back({0, {{u'0', u'9'}}, {}});
    Stemmers.push_back({2, {{u'а', u'я'}}, {u'ё'}});
    Stemmers.push_back({1, {{u'a', u'z'}}, {}});
    Stemmers.push_back({3, {{u'a', u'z'}}, {u'ẞ', u'ä', u'ü', u'ö'}});
    Stemmers.push_back({4, {{u'a', u'z'}}, {u'é', u'â', u'à', u'ê', u'è', u'ë', u'ç', u'î', u'ï', u'ô', u'û', u'ù', u'ü', u'ÿ'}});
}

static bool in_stemmer_range(const StemmerEntry& st, const u16s& w) {   // !Regex.IsMatch(word, "[^range]"), Build.cs:216
    for (char16_t c : w) {
        bool ok = false;
        for (auto& r : st.ranges) if (c >= r.first && c <= r.second) ok = true;
        for (char16_t s : st.singles) if (c == s) ok = true;
        if (!ok) return false;
    }
    return true;
}

static u16s hex_upper(uint32_t v) {   // String.Format("{0:X}", i), Index.cs:196
    char buf[16];
    snprintf(buf, sizeof buf, "%X", v);
    u16s r;
    for (char* p = buf; *p; p++) r.push_back((char16_t)*p);
    return r;
}

// GetWordCodes, Build.cs:175-247.  Returns {} both for "null" (empty word) and for stop words;
// callers on the path treat the two identically (Build.cs:306, Search.cs:226-227 throws->caught).
std::vector<u16s> Index::GetWordCodes(const u16s& word) const {
    std::vector<u16s> l;
    if (word.empty()) return l;
    if (word[0] <= u'9' && word[0] >= u'0') { l.push_back(word); return l; }
    u16s stemmed = word;
    int nG = 0;
    int nVoc = 0;
    if (stopWords.count(word)) return l;
    u16s firststemmed;
    l.push_back(word);
    for (const auto& voc : vocs) {
        // short-circuit chain of Build.cs:195 (assignments happen only while earlier clauses hold)
        bool hit = false;
        if (voc && word[0] >= voc->begin && word[0] <= voc->end) {
            stemmed = voc->stem(word);
            nG = voc->search(stemmed);
            hit = nG != 0;
        }
        if (hit) l.push_back(u"#" + hex_upper(((uint32_t)nVoc << 24) | ((uint32_t)nG & 0xFFFFFF)));
        if (firststemmed.empty()) firststemmed = stemmed;
        nVoc++;
    }
    if (nG == 0) {
        stemmed = firststemmed;
        if (vocs.empty()) {
            for (const auto& st : Stemmers) {
                if (in_stemmer_range(st, word)) {
                    if (st.kind == 1) stemmed = stem_english(word);
                    else if (st.kind == 2) stemmed = stem_russian(word);
                    break;
                }
            }
        }
        if (!stemmed.empty() && stemmed != word) l.push_back(u"$" + stemmed);
    }
    return l;
}

// =====================================================================================
// Builder  (Build.cs:258-437) and IndexTask (Build.cs:445-564)
// =====================================================================================
void Index::AddDoc(const u16s& sourceid, const u16s& name) {   // Build.cs:348-356
    pages.emplace_back(sourceid + u":" + name, maxCoord);
}
void Index::Add(const u16s& code, u64 c) {   // Build.cs:324-341 (chunk flush not modelled: canonical single chunk)
    maxCoord = c;
    builder[code].add(c);
    totalPostings++;
}
void Index::AddWord(const u16s& word, u64 c) {   // Build.cs:302-316 (code.suff is always null)
    for (const auto& code : GetWordCodes(word)) Add(code, c);
}
void Index::EndPage(const u16s& id, u64 c) { pages.emplace_back(u":" + id, c); }   // Build.cs:364-367

// StringReader.ReadLine: splits on \n, \r, \r\n; returns false at end.
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

// .NET Regex \b uses RegexCharClass.IsWordChar: \w plus U+200C / U+200D.
static inline bool is_bw(char16_t c) { return is_w(c) || c == 0x200C || c == 0x200D; }

// One iteration of the page loop of IndexTask, Build.cs:477-536.
void Index::IndexPage(const u16s& id, const u16s& text) {
    u16s c = to_lower(text);                                       // Build.cs:482
    if (c.empty()) return;                                         // Build.cs:483
    if (id == u"0") {                                              // Build.cs:485-524 header page
        size_t pos = 0;
        u16s line;
        while (read_line(text, pos, line)) {
            line = to_lower(line);                                 // Build.cs:498
            // Split('='), Build.cs:499
            std::vector<u16s> fields;
            {
                size_t st = 0;
                for (size_t i = 0; i <= line.size(); i++)
                    if (i == line.size() || line[i] == u'=') { fields.push_back(line.substr(st, i - st)); st = i + 1; }
            }
            if ((int)fields[0].size() >= MIN_WORD_LENGTH) {
                if (fields.size() < 2) return;   // IndexOutOfRangeException -> caught at Build.cs:537: page abandoned, no EndPage
                const u16s& v = fields[1];
                // Regex.Split(v, "\\b"), Build.cs:502: pieces are the maximal runs of equal word-ness
                // (plus empty strings at either end, which only matter through dc += Length).
                int dc = (int)fields[0].size() + 1;                // Build.cs:503
                size_t i = 0;
                while (i < v.size()) {
                    size_t j = i + 1;
                    bool w = is_bw(v[i]);
                    while (j < v.size() && is_bw(v[j]) == w) j++;
                    u16s piece = v.substr(i, j - i);
                    if (is_w(piece[0])) {                          // Build.cs:506
                        Add(u"&" + fields[0], (u64)(uint32_t)(coord + (u64)(dc - 1)));   // Build.cs:509 ((uint) cast)
                        AddWord(to_lower(piece), coord + (u64)dc);                      // Build.cs:510
                    }
                    dc += (int)piece.size();                       // Build.cs:512
                    i = j;
                }
            }
            coord += (u64)line.size() + 1;                         // Build.cs:515
        }
        EndPage(id, coord);                                        // Build.cs:521
        return;
    }
    // Regex.Matches(c, @"\p{L}+|\p{N}+"), Build.cs:526-530
    size_t n = c.size();
    size_t i = 0;
    while (i < n) {
        int k = is_L(c[i]) ? 1 : (is_N(c[i]) ? 2 : 0);
        if (!k) { i++; continue; }
        size_t j = i + 1;
        if (k == 1) while (j < n && is_L(c[j])) j++;
        else while (j < n && is_N(c[j])) j++;
        size_t len = j - i;
        if ((int)len >= MIN_WORD_LENGTH && (int)len <= MAX_WORD_LENGTH)
            AddWord(c.substr(i, len), coord + (uint32_t)i);        // Build.cs:529
        i = j;
    }
    coord += (uint32_t)n;                                          // Build.cs:531
    EndPage(id, coord);                                            // Build.cs:532
}

// BinaryWriter helpers
static void w_u64(std::vector<uint8_t>& o, u64 v) { for (int i = 0; i < 8; i++) o.push_back((uint8_t)(v >> (8 * i))); }
static void w_i32(std::vector<uint8_t>& o, int32_t v) { for (int i = 0; i < 4; i++) o.push_back((uint8_t)((uint32_t)v >> (8 * i))); }
static void w_str(std::vector<uint8_t>& o, const u16s& s) {
    std::string u = utf16_to_utf8(s);
    uint32_t n = (uint32_t)u.size();
    while (n >= 0x80) { o.push_back((uint8_t)(n | 0x80)); n >>= 7; }
    o.push_back((uint8_t)n);
    o.insert(o.end(), u.begin(), u.end());
}

// Builder.Save (Build.cs:370-404) -> MergeIndexes with one file (Index.cs:695-752; Shift(0) and the
// re-encode of MergeFiles:665-674 are identities) -> IndexPageList.Save (Build.cs:99-118) -> Load.
void Index::Finish() {
    index_bytes.clear();
    list_bytes.clear();
    w_u64(index_bytes, maxCoord);                                  // Build.cs:377
    for (const auto& kv : builder) {                               // Build.cs:379-384
        w_str(index_bytes, kv.first);
        const auto& self = kv.second.seq.self;
        w_i32(index_bytes, (int32_t)self.size());                  // IndexSequence.Write, IndexSequence.cs:167-173
        for (uint16_t w : self) { index_bytes.push_back((uint8_t)w); index_bytes.push_back((uint8_t)(w >> 8)); }
    }
    // AddFromList (Build.cs:53-72) + Save (Build.cs:99-118)
    int doc = -1, lastSavedDoc = -1;
    std::vector<u16s> docs;
    for (const auto& d : pages) {
        if (d.first[0] != u':') {
            docs.push_back(d.first);
            doc = (int)docs.size() - 1;
        } else {
            w_u64(list_bytes, d.second);
            if (doc != lastSavedDoc) {
                lastSavedDoc = doc;
                w_str(list_bytes, doc >= 0 ? docs[doc] : u16s());
                w_u64(list_bytes, d.second);
            }
            w_str(list_bytes, d.first);
        }
    }
    Load(index_bytes.data(), index_bytes.size(), list_bytes.data(), list_bytes.size());
}

struct Rd {
    const uint8_t* p; size_t n; size_t i = 0; bool eof = false;
    u64 r_u64() { if (i + 8 > n) { eof = true; return 0; } u64 v = 0; for (int k = 0; k < 8; k++) v |= (u64)p[i + k] << (8 * k); i += 8; return v; }
    int32_t r_i32() { if (i + 4 > n) { eof = true; return 0; } uint32_t v = 0; for (int k = 0; k < 4; k++) v |= (uint32_t)p[i + k] << (8 * k); i += 4; return (int32_t)v; }
    u16s r_str() {
        uint32_t len = 0; int sh = 0;
        while (true) {
            if (i >= n) { eof = true; return u16s(); }
            uint8_t b = p[i++]; len |= (uint32_t)(b & 0x7F) << sh; sh += 7;
            if (!(b & 0x80)) break;
        }
        if (i + len > n) { eof = true; return u16s(); }
        u16s s = utf8_to_utf16(p + i, len); i += len; return s;
    }
};

// Index.Load, Index.cs:312-380 (InMemory layout) + IndexPageList.Load, Build.cs:119-148
bool Index::Load(const uint8_t* ib, size_t in, const uint8_t* lb, size_t ln) {
    CanSearch = false;
    self.clear();
    PagesList.clear();
    DocsList.clear();
    Rd r{ib, in};
    MaxCoord = r.r_u64();
    while (!r.eof) {
        u16s s = r.r_str();
        if (r.eof) break;
        int32_t n = r.r_i32();
        if (r.eof) break;
        Seq q;
        if (r.i + (size_t)n * 2 > r.n) break;
        q.self.resize(n);
        for (int32_t k = 0; k < n; k++) q.self[k] = (uint16_t)(r.p[r.i + 2 * k] | (r.p[r.i + 2 * k + 1] << 8));
        r.i += (size_t)n * 2;
        self.emplace(s, std::move(q));
    }
    Rd l{lb, ln};
    int doc = -1;
    while (true) {
        u64 n = l.r_u64();
        if (l.eof) break;
        u16s s = l.r_str();
        if (l.eof) break;
        if (s.empty()) break;   // s[0] on empty string throws -> loop ends (Build.cs:141)
        if (s[0] != u':') { DocsList.push_back(s); doc = (int)DocsList.size() - 1; }
        else { DocPage p; p.doc = doc; p.id = s.substr(1); PagesList.emplace_back(n, p); }
    }
    CanSearch = true;
    return true;
}

// =====================================================================================
// Search  (Search.cs)
// =====================================================================================
// GetLikeWords, Search.cs:160-167: '_' -> ".*", unanchored Regex.IsMatch over keys whose first
// char is a letter, first MAX_LIKE_WORDS=100 in key order.  The pattern only contains \w
// characters and ".*", so matching == the literal segments occur in order.
std::vector<u16s> Index::GetLikeWords(const u16s& word) const {
    if (word.find(u'_') == u16s::npos || !bKeepForms) return {word};
    if (word.size() < 2) return {};
    std::vector<u16s> segs;
    {
        size_t st = 0;
        for (size_t i = 0; i <= word.size(); i++)
            if (i == word.size() || word[i] == u'_') { if (i > st) segs.push_back(word.substr(st, i - st)); st = i + 1; }
    }
    std::vector<u16s> out;
    for (const auto& kv : self) {
        const u16s& key = kv.first;
        if (key.empty() || !is_L(key[0])) continue;
        size_t pos = 0;
        bool ok = true;
        for (const auto& sg : segs) {
            size_t f = key.find(sg, pos);
            if (f == u16s::npos) { ok = false; break; }
            pos = f + sg.size();
        }
        if (ok) { out.push_back(key); if (out.size() >= 100) break; }
    }
    return out;
}

// SearchWord, Search.cs:192-260 (InMemory == false semantics: this[code] yields a fresh R == 0 copy)
Seq Index::SearchWord(const u16s& word_in) const {
    u16s word = word_in;
    bool bExact = false;
    if (to_upper(word) == word && bKeepForms) bExact = true;
    word = to_lower(word);
    bool have_total = false;
    Seq total;
    std::vector<u16s> words = {word};
    if (word.find(u'_') != u16s::npos) {
        if (bKeepForms) { bExact = true; words = GetLikeWords(word); }
        else return Seq();
    }
    for (const auto& wword : words) {
        std::vector<u16s> codes = GetWordCodes(wword);
        if (codes.empty()) continue;
        std::vector<u16s> selfcodes, knowcodes;
        for (const auto& s : codes) {
            if (is_w(s[0])) { if (std::find(selfcodes.begin(), selfcodes.end(), s) == selfcodes.end()) selfcodes.push_back(s); }
        }
        for (const auto& s : codes)   // codes.Except(selfcodes): distinct, order preserved
            if (std::find(selfcodes.begin(), selfcodes.end(), s) == selfcodes.end() &&
                std::find(knowcodes.begin(), knowcodes.end(), s) == knowcodes.end())
                knowcodes.push_back(s);
        std::vector<u16s> use;
        if (!bExact && !knowcodes.empty()) use = knowcodes;
        else if (!selfcodes.empty()) {
            // selfcodes is a lazy Where over `codes` (duplicates kept), Take(1) == first self code
            for (const auto& s : codes) if (is_w(s[0])) { use.push_back(s); break; }
        }
        for (const auto& code : use) {
            auto it = self.find(code);
            if (it == self.end()) continue;
            Seq res = it->second;
            res.R = 0;
            if (!have_total) { total = res; have_total = true; }
            else total = op_add(total, res);
        }
    }
    if (!have_total) total = Seq();
    if (bExact) total.R = -1;
    return total;
}

// SearchField, Search.cs:126-155
Seq Index::SearchField(const u16s& field, const u16s& value) const {
    auto it = self.find(u"&" + to_lower(field));
    if (it == self.end()) return Seq();   // KeyNotFoundException -> catch -> empty
    Seq seq = it->second;
    Seq newseq = SearchWord(to_lower(value));
    seq.R = -1;
    return op_mul(seq, newseq);
}

// ResultDocPage.rank, Search.cs:99-111
float ResultPage::rank() const {
    float bonus = 0;
    if (pos.size() > 1)
        for (size_t q = 1; q < pos.size(); q++) bonus += (float)(30 / std::max(5, pos[q] - pos[q - 1]));
    return 1 + bonus + (float)std::log((double)pos.size());
}

// PrepareSearchResult, Search.cs:365-420 with IndexPageList.GetPage, Build.cs:74-97 (the cached
// enumerator restarts when coord moves backwards, so the result is the first page with end > coord
// and page.coord = previous page's end (0 for the first page)).
SearchResult Index::PrepareSearchResult(const Seq& res, const std::vector<u16s>& filter) const {
    SearchResult result;
    int prevp = -1, prevd = -2;
    int lastDoc = -1;         // index into docs_all
    std::vector<ResultDoc> docs_all;   // every ResultDocument created (members of foundDocs are flagged)
    std::vector<char> in_found;
    size_t cur = 0;
    u64 prevc = 0;
    std::unordered_set<u16s> found_names;
    SeqEnum e(res);
    while (e.MoveNext()) {
        u64 coord = e.Current();
        // GetPage
        if (!(cur < PagesList.size() && PagesList[cur].first > coord && prevc <= coord)) {
            if (cur < PagesList.size() && PagesList[cur].first > coord) { cur = 0; prevc = 0; }
            while (cur < PagesList.size() && !(PagesList[cur].first > coord)) { prevc = PagesList[cur].first; cur++; }
        }
        if (cur >= PagesList.size()) break;   // new DocPage() with doc == null -> NullReferenceException -> Search's catch
        int p = (int)cur;
        u64 pcoord = prevc;
        const DocPage& pg = PagesList[p].second;
        if (p != prevp) {
            ResultPage rp;
            rp.page_index = p;
            rp.id = pg.id;
            rp.pos.push_back((int32_t)(coord - pcoord));
            result.foundPages.push_back(rp);
            if (prevd != pg.doc) {
                ResultDoc d;
                d.doc = pg.doc;
                d.name = pg.doc >= 0 ? DocsList[pg.doc] : u16s();
                bool added = false;
                if ((int)found_names.size() < MAX_FOUND_DOCS) {
                    bool matched = filter.empty();
                    // doc-name regex filters (Search.cs:389-395) are host-side and not on the measured path;
                    // the oracle supports literal substring filters only.
                    for (const auto& f : filter) if (d.name.find(f) != u16s::npos) { matched = true; break; }
                    if (matched && !found_names.count(d.name)) { found_names.insert(d.name); added = true; }
                }
                docs_all.push_back(d);
                in_found.push_back(added);
                lastDoc = (int)docs_all.size() - 1;
            }
            docs_all[lastDoc].pages.push_back((int)result.foundPages.size() - 1);
            prevp = p;
            prevd = pg.doc;
        } else {
            result.foundPages.back().pos.push_back((int32_t)(coord - pcoord));
        }
        if ((int)result.foundPages.size() > MAX_FOUND_PAGES) break;
    }
    for (size_t i = 0; i < docs_all.size(); i++) if (in_found[i]) result.foundDocs.push_back(docs_all[i]);
    return result;
}

namespace {

struct SearchSequence {   // Search.cs:280-317
    u16s Word;
    int kind;        // 0 = SearchWord, 1 = SearchField
    u16s field;
    int Dist = 0;
    WordInfo info;
};

struct Parser {   // C# precedence: '*' binds tighter than '+'; parentheses; variables are U+E000+i
    const Index* ix;
    std::vector<SearchSequence>* seq;
    const u16s* s;
    size_t i = 0;
    bool err = false;
    void ws() { while (i < s->size() && (*s)[i] == u' ') i++; }
    Seq primary() {
        ws();
        if (i >= s->size()) { err = true; return Seq(); }
        char16_t c = (*s)[i];
        if (c == u'(') {
            i++;
            Seq v = expr();
            ws();
            if (i >= s->size() || (*s)[i] != u')') { err = true; return Seq(); }
            i++;
            return v;
        }
        if (c >= 0xE000 && c < 0xE000 + seq->size()) {
            i++;
            SearchSequence& q = (*seq)[c - 0xE000];
            // SearchSequence.d(), Search.cs:302-312
            Seq res = q.kind == 0 ? ix->SearchWord(q.Word) : ix->SearchField(q.field, q.Word);
            q.info.nFound = (int)res.count_values();
            res.R = (res.R < 0 ? -(int)q.Word.size() - 4 : q.Dist + (int)q.Word.size());
            return res;
        }
        err = true;
        return Seq();
    }
    Seq term() {
        Seq a = primary();
        while (!err) {
            ws();
            if (i < s->size() && (*s)[i] == u'*') { i++; Seq b = primary(); if (!err) a = op_mul(a, b); }
            else break;
        }
        return a;
    }
    Seq expr() {
        Seq a = term();
        while (!err) {
            ws();
            if (i < s->size() && (*s)[i] == u'+') { i++; Seq b = term(); if (!err) a = op_add(a, b); }
            else break;
        }
        return a;
    }
    Seq parse() {
        Seq v = expr();
        ws();
        if (i != s->size()) err = true;
        return v;
    }
};

static inline bool bnd(const u16s& s, size_t i) {   // \b at position i
    bool a = i > 0 && is_bw(s[i - 1]);
    bool b = i < s.size() && is_bw(s[i]);
    return a != b;
}

// PrepareSearchRequest, Search.cs:319-363.  Variables are emitted as single chars U+E000+index
// instead of "A.d()" (the reference names them 'A'+index, Search.cs:357).
u16s PrepareSearchRequest(const Index& ix, u16s req, std::vector<SearchSequence>& funcs, u16s& fields, int kind,
                          const u16s& fieldname, bool keepshort) {
    fields.clear();
    // Search.cs:325  [^\w(){}=~?|"]|_+  -> " "
    {
        u16s o;
        for (size_t i = 0; i < req.size();) {
            char16_t c = req[i];
            if (c == u'_') { while (i < req.size() && req[i] == u'_') i++; o.push_back(u' '); continue; }
            bool keep = is_w(c) || c == u'(' || c == u')' || c == u'{' || c == u'}' || c == u'=' || c == u'~' || c == u'?' ||
                        c == u'|' || c == u'"';
            o.push_back(keep ? c : u' ');
            i++;
        }
        req = o;
    }
    // Search.cs:328-333  {*(\w+)[ ]*=([\w|() ]+)}
    u16s f;
    {
        u16s o;
        size_t i = 0;
        while (i < req.size()) {
            size_t j = i;
            while (j < req.size() && req[j] == u'{') j++;
            size_t k = j;
            while (k < req.size() && is_w(req[k])) k++;
            bool ok = k > j;
            size_t m = k;
            if (ok) { while (m < req.size() && req[m] == u' ') m++; ok = m < req.size() && req[m] == u'='; }
            size_t n = m + 1;
            if (ok) {
                while (n < req.size() && (is_w(req[n]) || req[n] == u'|' || req[n] == u'(' || req[n] == u')' || req[n] == u' ')) n++;
                ok = n > m + 1 && n < req.size() && req[n] == u'}';
            }
            if (ok) {
                u16s dummy;
                u16s l = PrepareSearchRequest(ix, req.substr(m + 1, n - (m + 1)), funcs, dummy, 1, req.substr(j, k - j), true);
                f += (f.empty() ? u"" : u"*");
                f += u"(" + l + u")";
                i = n + 1;
            } else {
                o.push_back(req[i]);
                i++;
            }
        }
        req = o;
    }
    fields = f;
    // Search.cs:337  {.*} -> ""
    {
        size_t a = req.find(u'{');
        size_t b = req.rfind(u'}');
        if (a != u16s::npos && b != u16s::npos && b > a) req.erase(a, b - a + 1);
    }
    for (auto& c : req) if (c == u'?') c = u'_';   // Search.cs:340
    // Search.cs:343-344  \b\w{1,2}\b -> " "
    if (!keepshort) {
        u16s o;
        for (size_t i = 0; i < req.size();) {
            if (is_w(req[i])) {
                size_t j = i;
                while (j < req.size() && is_w(req[j])) j++;
                bool whole = bnd(req, i) && bnd(req, j);
                if (j - i <= 2 && whole) o.push_back(u' ');
                else o.append(req, i, j - i);
                i = j;
            } else { o.push_back(req[i]); i++; }
        }
        req = o;
    }
    // Search.cs:346  \b{stop}\b -> ""
    for (const auto& st : ix.stopWords) {
        if (st.empty()) continue;
        u16s o;
        size_t i = 0;
        while (i < req.size()) {
            if (req.compare(i, st.size(), st) == 0 && bnd(req, i) && bnd(req, i + st.size())) i += st.size();
            else { o.push_back(req[i]); i++; }
        }
        req = o;
    }
    // Search.cs:349  "(.*)" -> (UPPER)
    {
        size_t a = req.find(u'"');
        size_t b = req.rfind(u'"');
        if (a != u16s::npos && b != u16s::npos && b > a)
            req = req.substr(0, a) + u"(" + to_upper(req.substr(a + 1, b - a - 1)) + u")" + req.substr(b + 1);
    }
    for (auto& c : req) if (c == u'|') c = u'+';   // Search.cs:351
    // Search.cs:353  (\b|\))(\s+)(\b|\() : the whitespace run becomes "*"
    {
        u16s o;
        size_t i = 0;
        while (i < req.size()) {
            size_t st = i;
            bool lead_paren = false;
            bool ok = false;
            size_t ws0 = i;
            if (bnd(req, i) && req[i] == u' ') { ok = true; ws0 = i; }
            else if (req[i] == u')' && i + 1 < req.size() && req[i + 1] == u' ') { ok = true; lead_paren = true; ws0 = i + 1; }
            if (ok) {
                size_t e = ws0;
                while (e < req.size() && req[e] == u' ') e++;
                bool tail_paren = false;
                bool tail_ok = false;
                if (e < req.size() && bnd(req, e)) tail_ok = true;
                else if (e < req.size() && req[e] == u'(') { tail_ok = true; tail_paren = true; }
                if (tail_ok) {
                    if (lead_paren) o.push_back(u')');
                    o.push_back(u'*');
                    if (tail_paren) { o.push_back(u'('); e++; }
                    i = e;
                    continue;
                }
            }
            o.push_back(req[st]);
            i = st + 1;
        }
        req = o;
    }
    // Search.cs:357  \b(\w+)\b -> variable
    {
        u16s o;
        for (size_t i = 0; i < req.size();) {
            if (is_w(req[i])) {
                size_t j = i;
                while (j < req.size() && is_w(req[j])) j++;
                SearchSequence q;
                q.Word = req.substr(i, j - i);
                q.kind = kind;
                q.field = fieldname;
                q.info.word = q.Word;
                funcs.push_back(q);
                o.push_back((char16_t)(0xE000 + funcs.size() - 1));
                i = j;
            } else { o.push_back(req[i]); i++; }
        }
        req = o;
    }
    return req;
}

}  // namespace

// Index.Search, Search.cs:437-617 (snippet/headers block 560-595 is host-side text work, not restated)
SearchResult Index::Search(const u16s& req_in, bool has_opt, int dist) const {
    SearchResult err;
    if (!CanSearch) { err.success = false; err.error = "Index is not built"; return err; }
    u16s req = to_lower(req_in);   // Search.cs:453
    // -filter: (Search.cs:456-466)  \B-filter:([\w\*\?\\.()+{}/]+,?)+
    std::vector<u16s> filter;
    {
        const u16s tag = u"-filter:";
        size_t i = 0;
        u16s o;
        auto fchar = [](char16_t c) {
            return is_w(c) || c == u'*' || c == u'?' || c == u'\\' || c == u'.' || c == u'(' || c == u')' || c == u'+' || c == u'{' ||
                   c == u'}' || c == u'/';
        };
        bool first = true;
        while (i < req.size()) {
            if (req.compare(i, tag.size(), tag) == 0 && !bnd(req, i)) {
                size_t j = i + tag.size();
                std::vector<u16s> caps;
                while (j < req.size() && fchar(req[j])) {
                    size_t k = j;
                    while (k < req.size() && fchar(req[k])) k++;
                    u16s cap = req.substr(j, k - j);
                    if (k < req.size() && req[k] == u',') k++;
                    caps.push_back(cap);
                    j = k;
                }
                if (!caps.empty()) {
                    if (first) { filter = caps; first = false; }   // Regex.Match: first match only feeds `filter`
                    o.push_back(u' ');
                    i = j;
                    continue;
                }
            }
            o.push_back(req[i]);
            i++;
        }
        req = o;
    }
    u16s fields;
    std::vector<SearchSequence> seq;
    req = PrepareSearchRequest(*this, req, seq, fields, 0, u16s(), false);   // Search.cs:470
    int R = 255;
    if (has_opt) R = dist;                                                  // Search.cs:472-473
    for (auto& s : seq) s.Dist = R;                                         // Search.cs:492
    bool have_res = false, have_resf = false;
    Seq res, resf;
    if (!req.empty()) {                                                     // Search.cs:497-509
        Parser p{this, &seq, &req};
        res = p.parse();
        if (p.err) { err.success = false; err.error = "Syntax Error in search request"; return err; }
        have_res = true;
    }
    if (!fields.empty()) {                                                  // Search.cs:525-537
        Parser p{this, &seq, &fields};
        resf = p.parse();
        if (p.err) { err.success = false; err.error = "Syntax Error in search request"; return err; }
        have_resf = true;
    }
    if (!have_res && have_resf) { res = resf; have_res = true; }            // Search.cs:538
    if (!have_res) return SearchResult();                                   // Search.cs:615-616
    SearchResult result = PrepareSearchResult(res, filter);                 // Search.cs:543
    if (have_resf) {                                                        // Search.cs:544-547, 423-428
        SearchResult rf = PrepareSearchResult(resf, {});
        std::unordered_set<u16s> names;
        for (auto& d : rf.foundDocs) names.insert(d.name);
        std::vector<ResultDoc> keep;
        for (auto& d : result.foundDocs) if (names.count(d.name)) keep.push_back(d);
        result.foundDocs = keep;
    }
    for (auto& doc : result.foundDocs) {                                    // Search.cs:552-557
        double sum = 0;   // Enumerable.Sum(float selector) accumulates in double
        for (int pi : doc.pages) sum += (double)result.foundPages[pi].rank();
        doc.rank = (float)sum;
        doc.rank = 1 + (float)std::log((double)doc.rank);
        if (!doc.pages.empty() && result.foundPages[doc.pages[0]].id == u"0") doc.rank *= 10.0f;   // DOC_RANK_MULTIPLY, Index.cs:115
    }
    std::stable_sort(result.foundDocs.begin(), result.foundDocs.end(),
                     [](const ResultDoc& a, const ResultDoc& b) { return a.rank < b.rank; });   // Search.cs:599
    for (auto& s : seq) result.words.push_back(s.info);                     // Search.cs:601
    return result;
}

}  // namespace orc
