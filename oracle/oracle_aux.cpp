// ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle_core.h).
// Dictionary-side queries over the key table and the .voc builders (row 4), restated from
// Docodo.NET/Search.cs:160-188, Index.cs:46-89,274-307 and Dict.cs:109-296.
#include <algorithm>
#include <stdexcept>
#include <cstring>

#include "oracle_core.h"

namespace orc {

// Utils.Levenshtein, Index.cs:46-89 (full matrix, unit costs)
static int levenshtein(const u16s& s, const u16s& t) {
    const int n = (int)s.size(), m = (int)t.size();
    if (n == 0) return m;
    if (m == 0) return n;
    std::vector<int> d((size_t)(n + 1) * (m + 1));
    auto D = [&](int i, int j) -> int& { return d[(size_t)i * (m + 1) + j]; };
    for (int i = 0; i <= n; i++) D(i, 0) = i;
    for (int j = 0; j <= m; j++) D(0, j) = j;
    for (int i = 1; i <= n; i++)
        for (int j = 1; j <= m; j++) {
            const int cost = t[j - 1] == s[i - 1] ? 0 : 1;
            D(i, j) = std::min(std::min(D(i - 1, j) + 1, D(i, j - 1) + 1), D(i - 1, j - 1) + cost);
        }
    return D(n, m);
}

// GetCloseWords, Search.cs:169-174: self.Keys.OrderBy(Levenshtein) (stable: ties stay in key order).Take(10)
std::vector<u16s> Index::GetCloseWords(const u16s& word) const {
    std::vector<std::pair<int, const u16s*>> v;
    for (const auto& kv : self) v.emplace_back(levenshtein(kv.first, word), &kv.first);
    std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    std::vector<u16s> out;
    for (size_t i = 0; i < v.size() && i < 10; i++) out.push_back(*v[i].second);
    return out;
}

// GetSuggessions, Search.cs:176-188
std::vector<u16s> Index::GetSuggessions(const u16s& req, int n) const {
    if (req.size() < 2) return {};
    // Regex.Split(req, @"\b").Last(s => s.Length > 0).ToLower(): the pieces are the maximal stretches between word boundaries
    auto bw = [](char16_t c) { return is_w(c) || c == 0x200C || c == 0x200D; };
    size_t st = 0;
    u16s last;
    for (size_t i = 1; i <= req.size(); i++) {
        const bool boundary = i == req.size() || (bw(req[i - 1]) != bw(req[i]));
        if (boundary) { if (i > st) last = req.substr(st, i - st); st = i; }
    }
    u16s lastword = to_lower(last);
    if (lastword.size() < 2) return {};
    std::vector<std::pair<long long, const u16s*>> v;
    for (const auto& kv : self) {
        const u16s& k = kv.first;
        if (k.empty() || k[0] < u'A') continue;
        if (k.size() <= lastword.size() || k.compare(0, lastword.size(), lastword) != 0) continue;
        v.emplace_back(-(long long)kv.second.self.size(), &k);   // OrderBy(-Value.Count): Count == number of ushorts
    }
    std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    std::vector<u16s> out;
    for (size_t i = 0; i < v.size() && (int)i < n; i++) out.push_back(v[i].second->substr(lastword.size()));
    return out;
}

// GetWordsGroup, Index.cs:274-282: first 20 words of the vocabulary group, joined by ","
u16s Index::GetWordsGroup(const u16s& hex) const {
    unsigned long long id = 0;
    for (char16_t c : hex) {
        int d = c >= u'0' && c <= u'9' ? c - u'0' : c >= u'A' && c <= u'F' ? c - u'A' + 10 : c >= u'a' && c <= u'f' ? c - u'a' + 10 : -1;
        if (d < 0) return u16s();
        id = id * 16 + (unsigned)d;
    }
    const size_t nV = (size_t)(id >> 24);
    if (nV >= vocs.size() || !vocs[nV]) return u16s();
    u16s out;
    int taken = 0;
    for (const auto& kv : vocs[nV]->ordered) {
        if (kv.second != (int32_t)(id & 0xFFFFFF)) continue;
        if (taken++) out += u",";
        out += kv.first;
        if (taken >= 20) break;
    }
    return out;
}

// CalcHistogram, Index.cs:284-307 for an InMemory index: the `top` keys with the longest sequences (Value.Count ushorts),
// ordered by -Count (stable: ties in key order); "#..." keys are shown as "(" + the words of the group + ")".  The reference
// adds to a Dictionary and stops at the first duplicate display key (the exception ends the loop, Index.cs:302-305).
std::vector<std::pair<u16s, int>> Index::CalcHistogram(size_t top) const {
    std::vector<std::pair<long long, const u16s*>> v;
    for (const auto& kv : self) v.emplace_back(-(long long)kv.second.self.size(), &kv.first);
    std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    std::vector<std::pair<u16s, int>> out;
    std::set<u16s> seen;
    for (size_t i = 0; i < v.size() && i < top; i++) {
        const u16s& key = *v[i].second;
        u16s shown = key;
        if (!key.empty() && key[0] == u'#') {
            const u16s g = GetWordsGroup(key.substr(1));
            if (g.empty()) break;   // Aggregate over an empty group throws InvalidOperationException: the loop ends (Index.cs:302-305)
            shown = u"(" + g + u")";
        }
        if (!seen.insert(shown).second) break;
        out.emplace_back(shown, (int)-v[i].first);
    }
    return out;
}

// ---- VocBuilder.AddWordsGroup, Dict.cs:120-190 ------------------------------------------------------------------
static const int32_t NOT_EXACT = 0x01000000, NUMBER_MASK = 0xFFFFFF;   // Dict.cs:15-16
void VocBuilder::AddWordsGroup(const std::vector<u16s>& grouplist) {
    auto stem = [&](const u16s& w) { return stemmer == 1 ? stem_english(w) : stemmer == 2 ? stem_russian(w) : w; };
    int32_t currNG = nG;
    bool hasmatch = false, found = false;
    std::set<int32_t> replaceGroups;
    for (const auto& word : grouplist) {
        const u16s stemme = stem(word);
        if (!hasmatch && std::find(grouplist.begin(), grouplist.end(), stemme) != grouplist.end()) hasmatch = true;
        auto it = dict.find(stemme);
        if (it != dict.end()) {
            int32_t newcurrNG = it->second;
            auto rp = replaces.find(newcurrNG);
            if (rp != replaces.end()) newcurrNG = rp->second;
            if ((currNG & NUMBER_MASK) != (newcurrNG & NUMBER_MASK)) {
                if (found) replaceGroups.insert(newcurrNG & NUMBER_MASK);
                else currNG = newcurrNG;
                found = true;
            }
        }
    }
    if ((currNG & NOT_EXACT) == 0) hasmatch = true;
    if (hasmatch) currNG &= ~NOT_EXACT;
    for (int32_t gr : replaceGroups) {
        auto rp = replaces.find(gr);
        if (rp != replaces.end()) { if (rp->second != currNG) throw std::runtime_error("Error duplicate replaces"); }
        else replaces.emplace(gr, currNG);
    }
    for (const auto& word : grouplist) {
        const u16s stemme = stem(word);
        auto it = dict.find(stemme);
        if (it == dict.end()) dict.emplace(stemme, currNG);
        else if (hasmatch && (it->second & NOT_EXACT) != 0) it->second = currNG & ~NOT_EXACT;
    }
    nG++;
}

// VocBuilder.build, Dict.cs:192-208: {BinaryWriter string, int32} in ordinal key order, group ids through `replaces`
std::vector<uint8_t> VocBuilder::build() const {
    std::vector<uint8_t> out;
    for (const auto& kv : dict) {
        const std::string k = utf16_to_utf8(kv.first);
        size_t n = k.size();
        while (n >= 0x80) { out.push_back((uint8_t)((n & 0x7F) | 0x80)); n >>= 7; }
        out.push_back((uint8_t)n);
        out.insert(out.end(), k.begin(), k.end());
        int32_t v = kv.second;
        auto rp = replaces.find(v);
        if (rp != replaces.end()) v = rp->second;
        uint8_t b[4];
        memcpy(b, &v, 4);
        out.insert(out.end(), b, b + 4);
    }
    return out;
}

}  // namespace orc
