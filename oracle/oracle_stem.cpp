// ORACLE -- TEST INFRASTRUCTURE ONLY (see oracle_core.h).
//
// Snowball stemmers restated from the published algorithm definitions
// (snowballstem.org "The English (Porter2) stemming algorithm" / "Russian stemming algorithm").
// The reference calls Iveonik.Stemmers.EnglishStemmer / RussianStemmer from the NuGet package
// StemmersNet.Standard 1.1.1 (Docodo.NET.csproj:38) -- a C# port of the ~2010 Snowball
// distribution that is NOT vendored in /root/reference.  Call sites: Dict.cs:65 (Vocab.Stem),
// Build.cs:222 (stemmer-only branch), Dict.cs:130,177 (voc builder).
//   English: pinned by Dict/en.voc (keys are stems produced by that very stemmer,
//            Dict.cs:260-296) -- see tests/test_oracle_stemmer.py.
//   Russian: PARITY UNPINNED (Dict/ru.voc is a missing blob; no test vector in the reference).
// The ~2010 definition is used on purpose: no "ogist"->"og" rule, no later exception-list
// additions, no Russian yo->ye normalisation (all introduced in Snowball >= 2.0).
#include "oracle_core.h"

namespace orc {
namespace {

// ------------------------------------------------------------------------------------------
// English (Porter2)
// ------------------------------------------------------------------------------------------
struct En {
    u16s s;
    size_t p1 = 0, p2 = 0;

    static bool vowel(char16_t c) { return c == 'a' || c == 'e' || c == 'i' || c == 'o' || c == 'u' || c == 'y'; }
    bool ends(const char* suf) const {
        size_t n = 0;
        while (suf[n]) n++;
        if (n > s.size()) return false;
        for (size_t i = 0; i < n; i++)
            if (s[s.size() - n + i] != (char16_t)(unsigned char)suf[i]) return false;
        return true;
    }
    // longest suffix among a list; returns index or -1, sets len
    int among(const char* const* list, int n, size_t& len) const {
        int best = -1;
        len = 0;
        for (int i = 0; i < n; i++) {
            size_t l = 0;
            while (list[i][l]) l++;
            if (l >= len && ends(list[i]) && (best < 0 || l > len)) { best = i; len = l; }
        }
        return best;
    }
    void replace_suffix(size_t len, const char* with) {
        s.resize(s.size() - len);
        for (const char* q = with; *q; q++) s.push_back((char16_t)(unsigned char)*q);
    }
    // "ends in a short syllable" evaluated for the prefix s[0..end)
    bool shortv(size_t end) const {
        if (end >= 3) {
            char16_t c = s[end - 1], v = s[end - 2], b = s[end - 3];
            if (!vowel(c) && c != 'w' && c != 'x' && c != 'Y' && vowel(v) && !vowel(b)) return true;
        }
        if (end == 2 && vowel(s[0]) && !vowel(s[1])) return true;
        return false;
    }
    bool has_vowel(size_t end) const {
        for (size_t i = 0; i < end; i++) if (vowel(s[i])) return true;
        return false;
    }

    void prelude() {
        if (!s.empty() && s[0] == '\'') s.erase(0, 1);
        if (!s.empty() && s[0] == 'y') s[0] = 'Y';
        for (size_t i = 1; i < s.size(); i++)
            if (s[i] == 'y' && vowel(s[i - 1])) s[i] = 'Y';
    }
    void mark_regions() {
        size_t n = s.size();
        p1 = p2 = n;
        size_t i = 0;
        auto starts = [&](const char* pre) {
            size_t l = 0;
            while (pre[l]) l++;
            if (l > n) return false;
            for (size_t k = 0; k < l; k++) if (s[k] != (char16_t)(unsigned char)pre[k]) return false;
            return true;
        };
        if (starts("gener") || starts("arsen")) i = 5;
        else if (starts("commun")) i = 6;
        else {
            while (i < n && !vowel(s[i])) i++;
            if (i >= n) return;
            while (i < n && vowel(s[i])) i++;
            if (i >= n) return;
            i++;
        }
        p1 = i;
        while (i < n && !vowel(s[i])) i++;
        if (i >= n) return;
        while (i < n && vowel(s[i])) i++;
        if (i >= n) return;
        p2 = i + 1;
    }
    void step_1a() {
        static const char* const apo[] = {"'s'", "'s", "'"};
        size_t len;
        if (among(apo, 3, len) >= 0) s.resize(s.size() - len);
        static const char* const l[] = {"sses", "ied", "ies", "s", "us", "ss"};
        int k = among(l, 6, len);
        switch (k) {
            case 0: replace_suffix(4, "ss"); break;
            case 1: case 2:
                if (s.size() - 3 > 1) replace_suffix(3, "i"); else replace_suffix(3, "ie");
                break;
            case 3:
                if (s.size() >= 2 && has_vowel(s.size() - 2)) s.resize(s.size() - 1);
                break;
            default: break;
        }
    }
    void step_1b() {
        static const char* const l[] = {"eed", "eedly", "ed", "edly", "ing", "ingly"};
        size_t len;
        int k = among(l, 6, len);
        if (k < 0) return;
        if (k <= 1) {
            if (s.size() - len >= p1) replace_suffix(len, "ee");
            return;
        }
        if (!has_vowel(s.size() - len)) return;
        s.resize(s.size() - len);
        if (ends("at") || ends("bl") || ends("iz")) { s.push_back('e'); return; }
        static const char* const dbl[] = {"bb", "dd", "ff", "gg", "mm", "nn", "pp", "rr", "tt"};
        size_t dl;
        if (among(dbl, 9, dl) >= 0) { s.resize(s.size() - 1); return; }
        if (s.size() == p1 && shortv(s.size())) s.push_back('e');
    }
    void step_1c() {
        size_t n = s.size();
        if (n >= 3 && (s[n - 1] == 'y' || s[n - 1] == 'Y') && !vowel(s[n - 2])) s[n - 1] = 'i';
    }
    void step_2() {
        static const char* const l[] = {"tional", "enci", "anci", "abli", "entli", "izer", "ization", "ational",
                                        "ation", "ator", "alism", "aliti", "alli", "fulness", "ousli", "ousness",
                                        "iveness", "iviti", "biliti", "bli", "ogi", "fulli", "lessli", "li"};
        static const char* const r[] = {"tion", "ence", "ance", "able", "ent", "ize", "ize", "ate",
                                        "ate", "ate", "al", "al", "al", "ful", "ous", "ous",
                                        "ive", "ive", "ble", "ble", "og", "ful", "less", ""};
        size_t len;
        int k = among(l, 24, len);
        if (k < 0 || s.size() - len < p1) return;
        if (k == 20) {   // ogi
            if (s.size() > len && s[s.size() - len - 1] == 'l') replace_suffix(len, "og");
            return;
        }
        if (k == 23) {   // li
            if (s.size() > len) {
                char16_t c = s[s.size() - len - 1];
                if (c == 'c' || c == 'd' || c == 'e' || c == 'g' || c == 'h' || c == 'k' || c == 'm' || c == 'n' ||
                    c == 'r' || c == 't')
                    s.resize(s.size() - 2);
            }
            return;
        }
        replace_suffix(len, r[k]);
    }
    void step_3() {
        static const char* const l[] = {"tional", "ational", "alize", "icate", "iciti", "ical", "ful", "ness", "ative"};
        static const char* const r[] = {"tion", "ate", "al", "ic", "ic", "ic", "", "", ""};
        size_t len;
        int k = among(l, 9, len);
        if (k < 0 || s.size() - len < p1) return;
        if (k == 8) {
            if (s.size() - len >= p2) s.resize(s.size() - len);
            return;
        }
        replace_suffix(len, r[k]);
    }
    void step_4() {
        static const char* const l[] = {"al", "ance", "ence", "er", "ic", "able", "ible", "ant", "ement", "ment",
                                        "ent", "ism", "ate", "iti", "ous", "ive", "ize", "ion"};
        size_t len;
        int k = among(l, 18, len);
        if (k < 0 || s.size() - len < p2) return;
        if (k == 17) {
            if (s.size() > len) {
                char16_t c = s[s.size() - len - 1];
                if (c == 's' || c == 't') s.resize(s.size() - len);
            }
            return;
        }
        s.resize(s.size() - len);
    }
    void step_5() {
        size_t n = s.size();
        if (n == 0) return;
        if (s[n - 1] == 'e') {
            if (n - 1 >= p2 || (n - 1 >= p1 && !shortv(n - 1))) s.resize(n - 1);
        } else if (s[n - 1] == 'l') {
            if (n - 1 >= p2 && n >= 2 && s[n - 2] == 'l') s.resize(n - 1);
        }
    }
    void postlude() {
        for (auto& c : s) if (c == 'Y') c = 'y';
    }
};

struct Exc { const char* from; const char* to; };
static const Exc kException1[] = {
    {"skis", "ski"}, {"skies", "sky"}, {"dying", "die"}, {"lying", "lie"}, {"tying", "tie"},
    {"idly", "idl"}, {"gently", "gentl"}, {"ugly", "ugli"}, {"early", "earli"}, {"only", "onli"},
    {"singly", "singl"}, {"sky", "sky"}, {"news", "news"}, {"howe", "howe"}, {"atlas", "atlas"},
    {"cosmos", "cosmos"}, {"bias", "bias"}, {"andes", "andes"}};
static const char* const kException2[] = {"inning", "outing", "canning", "herring", "earring", "proceed", "exceed", "succeed"};

static bool eq_ascii(const u16s& s, const char* a) {
    size_t i = 0;
    for (; a[i]; i++) if (i >= s.size() || s[i] != (char16_t)(unsigned char)a[i]) return false;
    return i == s.size();
}
static u16s from_ascii(const char* a) {
    u16s r;
    for (; *a; a++) r.push_back((char16_t)(unsigned char)*a);
    return r;
}

}  // namespace

u16s stem_english(const u16s& w) {
    for (const auto& e : kException1) if (eq_ascii(w, e.from)) return from_ascii(e.to);
    if (w.size() < 3) return w;
    En st;
    st.s = w;
    st.prelude();
    st.mark_regions();
    st.step_1a();
    bool exc2 = false;
    for (const char* e : kException2) if (eq_ascii(st.s, e)) exc2 = true;
    if (!exc2) {
        st.step_1b();
        st.step_1c();
        st.step_2();
        st.step_3();
        st.step_4();
        st.step_5();
    }
    st.postlude();
    return st.s;
}

// ------------------------------------------------------------------------------------------
// Russian (Snowball).  PARITY UNPINNED.
// ------------------------------------------------------------------------------------------
namespace {

struct Ru {
    u16s s;
    size_t pV = 0, p2 = 0;
    static bool vowel(char16_t c) {
        return c == u'а' || c == u'е' || c == u'и' || c == u'о' || c == u'у' || c == u'ы' || c == u'э' || c == u'ю' || c == u'я';
    }
    bool ends_at(const u16s& suf, size_t end) const {
        if (suf.size() > end) return false;
        return s.compare(end - suf.size(), suf.size(), suf) == 0;
    }
    // longest suffix of s[0..end) in list, restricted to start >= pV (the algorithm runs inside RV)
    int among(const std::vector<u16s>& list, size_t end, size_t& len) const {
        int best = -1;
        len = 0;
        for (size_t i = 0; i < list.size(); i++) {
            size_t l = list[i].size();
            if (end >= l && end - l >= pV && ends_at(list[i], end) && (best < 0 || l > len)) { best = (int)i; len = l; }
        }
        return best;
    }
    void mark_regions() {
        size_t n = s.size();
        pV = p2 = n;
        size_t i = 0;
        while (i < n && !vowel(s[i])) i++;
        if (i >= n) return;
        pV = i + 1;
        while (i < n && vowel(s[i])) i++;      // gopast non-v
        if (i >= n) return;
        i++;
        while (i < n && !vowel(s[i])) i++;     // gopast v
        if (i >= n) return;
        i++;
        while (i < n && vowel(s[i])) i++;      // gopast non-v
        if (i >= n) return;
        p2 = i + 1;
    }
};

static const std::vector<u16s> kPG1 = {u"в", u"вши", u"вшись"};                  // preceded by а or я
static const std::vector<u16s> kPG2 = {u"ив", u"ивши", u"ившись", u"ыв", u"ывши", u"ывшись"};
static const std::vector<u16s> kAdj = {u"ее", u"ие", u"ые", u"ое", u"ими", u"ыми", u"ей", u"ий", u"ый", u"ой", u"ем", u"им",
                                       u"ым", u"ом", u"его", u"ого", u"ему", u"ому", u"их", u"ых", u"ую", u"юю", u"ая", u"яя",
                                       u"ою", u"ею"};
static const std::vector<u16s> kPart1 = {u"ем", u"нн", u"вш", u"ющ", u"щ"};      // preceded by а or я
static const std::vector<u16s> kPart2 = {u"ивш", u"ывш", u"ующ"};
static const std::vector<u16s> kRefl = {u"ся", u"сь"};
static const std::vector<u16s> kVerb1 = {u"ла", u"на", u"ете", u"йте", u"ли", u"й", u"л", u"ем", u"н", u"ло", u"но", u"ет", u"ют",
                                         u"ны", u"ть", u"ешь", u"нно"};           // preceded by а or я
static const std::vector<u16s> kVerb2 = {u"ила", u"ыла", u"ена", u"ейте", u"уйте", u"ите", u"или", u"ыли", u"ей", u"уй", u"ил",
                                         u"ыл", u"им", u"ым", u"ен", u"ило", u"ыло", u"ено", u"ят", u"ует", u"уют", u"ит", u"ыт",
                                         u"ены", u"ить", u"ыть", u"ишь", u"ую", u"ю"};
static const std::vector<u16s> kNoun = {u"а", u"ев", u"ов", u"ие", u"ье", u"е", u"иями", u"ями", u"ами", u"еи", u"ии", u"и", u"ией",
                                        u"ей", u"ой", u"ий", u"й", u"иям", u"ям", u"ием", u"ем", u"ам", u"ом", u"о", u"у", u"ах",
                                        u"иях", u"ях", u"ы", u"ь", u"ию", u"ью", u"ю", u"ия", u"ья", u"я"};
static const std::vector<u16s> kDeriv = {u"ост", u"ость"};
static const std::vector<u16s> kSuper = {u"ейш", u"ейше"};

// Snowball `among` picks the longest suffix over the UNION of group-1 and group-2 endings; a
// group-1 ending then additionally needs а/я in front (which stays).  Returns new end or npos.
static size_t try_groups(const Ru& r, const std::vector<u16s>& g1, const std::vector<u16s>& g2, size_t end) {
    size_t l1, l2;
    int k1 = r.among(g1, end, l1), k2 = r.among(g2, end, l2);
    if (k1 < 0 && k2 < 0) return (size_t)-1;
    if (k2 >= 0 && (k1 < 0 || l2 >= l1)) return end - l2;
    // group 1 is the longest match
    size_t st = end - l1;
    if (st > r.pV - 0 && st >= 1 && st - 1 >= r.pV && (r.s[st - 1] == u'а' || r.s[st - 1] == u'я')) return st;
    return (size_t)-1;
}

}  // namespace

u16s stem_russian(const u16s& w) {
    Ru r;
    r.s = w;
    r.mark_regions();
    size_t end = r.s.size();
    if (r.pV >= end && r.pV == r.s.size()) {
        // RV empty: nothing can be removed
    }
    // Step 1
    size_t e = try_groups(r, kPG1, kPG2, end);
    if (e != (size_t)-1) {
        end = e;
    } else {
        size_t len;
        if (r.among(kRefl, end, len) >= 0) end -= len;
        bool done = false;
        // adjectival: adjective ending, optionally preceded by a participle ending
        if (r.among(kAdj, end, len) >= 0) {
            end -= len;
            size_t pe = try_groups(r, kPart1, kPart2, end);
            if (pe != (size_t)-1) end = pe;
            done = true;
        }
        if (!done) {
            size_t ve = try_groups(r, kVerb1, kVerb2, end);
            if (ve != (size_t)-1) { end = ve; done = true; }
        }
        if (!done) {
            if (r.among(kNoun, end, len) >= 0) end -= len;
        }
    }
    // Step 2: remove и
    if (end > r.pV && r.s[end - 1] == u'и') end--;
    // Step 3: derivational in R2
    {
        size_t len;
        Ru t = r;
        t.pV = r.pV;
        int k = t.among(kDeriv, end, len);
        if (k >= 0 && end - len >= r.p2) end -= len;
    }
    // Step 4: tidy up
    {
        size_t len;
        if (r.ends_at(u"нн", end) && end - 2 >= r.pV) {
            end -= 1;
        } else if (r.among(kSuper, end, len) >= 0) {
            end -= len;
            if (r.ends_at(u"нн", end) && end - 2 >= r.pV) end -= 1;
        } else if (end > r.pV && r.s[end - 1] == u'ь') {
            end -= 1;
        }
    }
    return r.s.substr(0, end);
}

}  // namespace orc
