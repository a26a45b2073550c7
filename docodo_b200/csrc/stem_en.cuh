// English Snowball (Porter2) stemmer, in place on a UTF-16 buffer, host+device.
//
// Replaces Iveonik.Stemmers.EnglishStemmer (NuGet StemmersNet.Standard 1.1.1, Docodo.NET.csproj:38) as called by
// Vocab.Stem (Dict.cs:65) and the stemmer-only branch of GetWordCodes (Build.cs:211-231).  Written from the
// published algorithm definition ("The English (Porter2) stemming algorithm", snowballstem.org), ~2010 revision,
// which is what StemmersNet ports: exception lists as published then, no "ogist" rule.
// Runs once per DISTINCT word (K4), never per token.  Output is never longer than the input.
#pragma once
#include "common.cuh"

namespace dcd {
namespace en {

DCD_HD bool vow(u16 c) { return c == 'a' || c == 'e' || c == 'i' || c == 'o' || c == 'u' || c == 'y'; }

template <int L>
DCD_HD bool ends(const u16* s, int n, const char (&suf)[L]) {
    const int l = L - 1;
    if (l > n) return false;
#pragma unroll
    for (int i = 0; i < l; i++)
        if (s[n - l + i] != (u16)(unsigned char)suf[i]) return false;
    return true;
}
template <int L>
DCD_HD bool whole(const u16* s, int n, const char (&w)[L]) { return n == L - 1 && ends(s, n, w); }
template <int L>
DCD_HD int put(u16* s, int n, const char (&t)[L]) {   // append literal at position n, return new length
#pragma unroll
    for (int i = 0; i < L - 1; i++) s[n + i] = (u16)(unsigned char)t[i];
    return n + L - 1;
}
DCD_HD bool has_vowel(const u16* s, int end) {
    for (int i = 0; i < end; i++)
        if (vow(s[i])) return true;
    return false;
}
// s[0..end) ends in a short syllable
DCD_HD bool short_syl(const u16* s, int end) {
    if (end >= 3) {
        u16 c = s[end - 1];
        if (!vow(c) && c != 'w' && c != 'x' && c != 'Y' && vow(s[end - 2]) && !vow(s[end - 3])) return true;
    }
    return end == 2 && vow(s[0]) && !vow(s[1]);
}

// Returns the new length.  `s` must hold at least n units; n may be anything up to MAX_KEY_UNITS.
DCD_HD int stem(u16* s, int n) {
    // exception1: whole-word special cases
    if (whole(s, n, "skis")) return 3;
    if (whole(s, n, "skies")) { s[2] = 'y'; return 3; }
    if (whole(s, n, "dying")) { s[1] = 'i'; s[2] = 'e'; return 3; }
    if (whole(s, n, "lying")) { s[1] = 'i'; s[2] = 'e'; return 3; }
    if (whole(s, n, "tying")) { s[1] = 'i'; s[2] = 'e'; return 3; }
    if (whole(s, n, "idly")) return 3;
    if (whole(s, n, "gently")) return 5;
    if (whole(s, n, "ugly")) { s[3] = 'i'; return 4; }
    if (whole(s, n, "early")) { s[4] = 'i'; return 5; }
    if (whole(s, n, "only")) { s[3] = 'i'; return 4; }
    if (whole(s, n, "singly")) return 5;
    if (whole(s, n, "sky") || whole(s, n, "news") || whole(s, n, "howe") || whole(s, n, "atlas") || whole(s, n, "cosmos") ||
        whole(s, n, "bias") || whole(s, n, "andes"))
        return n;
    if (n < 3) return n;

    // prelude
    if (s[0] == '\'') {
        for (int i = 1; i < n; i++) s[i - 1] = s[i];
        n--;
    }
    if (n > 0 && s[0] == 'y') s[0] = 'Y';
    for (int i = 1; i < n; i++)
        if (s[i] == 'y' && vow(s[i - 1])) s[i] = 'Y';

    // regions
    int p1 = n, p2 = n;
    {
        int i = 0;
        bool ok = true;
        if (n >= 5 && s[0] == 'g' && s[1] == 'e' && s[2] == 'n' && s[3] == 'e' && s[4] == 'r') i = 5;
        else if (n >= 5 && s[0] == 'a' && s[1] == 'r' && s[2] == 's' && s[3] == 'e' && s[4] == 'n') i = 5;
        else if (n >= 6 && s[0] == 'c' && s[1] == 'o' && s[2] == 'm' && s[3] == 'm' && s[4] == 'u' && s[5] == 'n') i = 6;
        else {
            while (i < n && !vow(s[i])) i++;
            while (i < n && vow(s[i])) i++;
            if (i >= n) ok = false; else i++;
        }
        if (ok) {
            p1 = i;
            while (i < n && !vow(s[i])) i++;
            while (i < n && vow(s[i])) i++;
            if (i < n) p2 = i + 1;
        }
    }

    // step 1a
    if (ends(s, n, "'s'")) n -= 3;
    else if (ends(s, n, "'s")) n -= 2;
    else if (ends(s, n, "'")) n -= 1;
    if (ends(s, n, "sses")) n -= 2;
    else if (ends(s, n, "ied") || ends(s, n, "ies")) n -= (n - 3 > 1) ? 2 : 1;
    else if (ends(s, n, "us") || ends(s, n, "ss")) {
    } else if (ends(s, n, "s")) {
        if (n >= 2 && has_vowel(s, n - 2)) n--;
    }

    // exception2
    bool stop = whole(s, n, "inning") || whole(s, n, "outing") || whole(s, n, "canning") || whole(s, n, "herring") ||
                whole(s, n, "earring") || whole(s, n, "proceed") || whole(s, n, "exceed") || whole(s, n, "succeed");
    if (!stop) {
        // step 1b
        {
            int l = 0;
            bool eed = false;
            if (ends(s, n, "eedly")) { l = 5; eed = true; }
            else if (ends(s, n, "ingly")) l = 5;
            else if (ends(s, n, "edly")) l = 4;
            else if (ends(s, n, "eed")) { l = 3; eed = true; }
            else if (ends(s, n, "ing")) l = 3;
            else if (ends(s, n, "ed")) l = 2;
            if (l) {
                if (eed) {
                    if (n - l >= p1) n = n - l + 2;   // -> "ee"
                } else if (has_vowel(s, n - l)) {
                    n -= l;
                    if (ends(s, n, "at") || ends(s, n, "bl") || ends(s, n, "iz")) s[n++] = 'e';
                    else if (n >= 2 && s[n - 1] == s[n - 2] &&
                             (s[n - 1] == 'b' || s[n - 1] == 'd' || s[n - 1] == 'f' || s[n - 1] == 'g' || s[n - 1] == 'm' ||
                              s[n - 1] == 'n' || s[n - 1] == 'p' || s[n - 1] == 'r' || s[n - 1] == 't'))
                        n--;
                    else if (n == p1 && short_syl(s, n)) s[n++] = 'e';
                }
            }
        }
        // step 1c
        if (n >= 3 && (s[n - 1] == 'y' || s[n - 1] == 'Y') && !vow(s[n - 2])) s[n - 1] = 'i';

        // step 2 (longest suffix first, then the R1 test)
        {
#define DCD_S2(suf, rep)                                      \
    if (!done && ends(s, n, suf)) {                            \
        done = true;                                           \
        if (n - (int)(sizeof(suf) - 1) >= p1) n = put(s, n - (int)(sizeof(suf) - 1), rep); \
    }
            bool done = false;
            DCD_S2("ization", "ize") DCD_S2("ational", "ate") DCD_S2("fulness", "ful") DCD_S2("ousness", "ous")
            DCD_S2("iveness", "ive") DCD_S2("tional", "tion") DCD_S2("biliti", "ble") DCD_S2("lessli", "less")
            DCD_S2("entli", "ent") DCD_S2("ation", "ate") DCD_S2("alism", "al") DCD_S2("aliti", "al")
            DCD_S2("ousli", "ous") DCD_S2("iviti", "ive") DCD_S2("fulli", "ful") DCD_S2("enci", "ence")
            DCD_S2("anci", "ance") DCD_S2("abli", "able") DCD_S2("izer", "ize") DCD_S2("ator", "ate")
            DCD_S2("alli", "al") DCD_S2("bli", "ble")
            if (!done && ends(s, n, "ogi")) {
                done = true;
                if (n - 3 >= p1 && n >= 4 && s[n - 4] == 'l') n -= 1;
            }
            if (!done && ends(s, n, "li")) {
                done = true;
                if (n - 2 >= p1 && n >= 3) {
                    u16 c = s[n - 3];
                    if (c == 'c' || c == 'd' || c == 'e' || c == 'g' || c == 'h' || c == 'k' || c == 'm' || c == 'n' || c == 'r' ||
                        c == 't')
                        n -= 2;
                }
            }
#undef DCD_S2
        }
        // step 3
        {
            bool done = false;
#define DCD_S3(suf, rep)                                      \
    if (!done && ends(s, n, suf)) {                            \
        done = true;                                           \
        if (n - (int)(sizeof(suf) - 1) >= p1) n = put(s, n - (int)(sizeof(suf) - 1), rep); \
    }
            DCD_S3("ational", "ate") DCD_S3("tional", "tion") DCD_S3("alize", "al") DCD_S3("icate", "ic")
            DCD_S3("iciti", "ic")
            if (!done && ends(s, n, "ative")) {
                done = true;
                if (n - 5 >= p1 && n - 5 >= p2) n -= 5;
            }
            DCD_S3("ical", "ic") DCD_S3("ness", "") DCD_S3("ful", "")
#undef DCD_S3
        }
        // step 4 (R2)
        {
            bool done = false;
#define DCD_S4(suf)                                            \
    if (!done && ends(s, n, suf)) {                            \
        done = true;                                           \
        if (n - (int)(sizeof(suf) - 1) >= p2) n -= (int)(sizeof(suf) - 1); \
    }
            DCD_S4("ement") DCD_S4("ance") DCD_S4("ence") DCD_S4("able") DCD_S4("ible") DCD_S4("ment")
            DCD_S4("ant") DCD_S4("ent") DCD_S4("ism") DCD_S4("ate") DCD_S4("iti") DCD_S4("ous") DCD_S4("ive") DCD_S4("ize")
            if (!done && ends(s, n, "ion")) {
                done = true;
                if (n - 3 >= p2 && n >= 4 && (s[n - 4] == 's' || s[n - 4] == 't')) n -= 3;
            }
            DCD_S4("al") DCD_S4("er") DCD_S4("ic")
#undef DCD_S4
        }
        // step 5
        if (n > 0 && s[n - 1] == 'e') {
            if (n - 1 >= p2 || (n - 1 >= p1 && !short_syl(s, n - 1))) n--;
        } else if (n > 1 && s[n - 1] == 'l') {
            if (n - 1 >= p2 && s[n - 2] == 'l') n--;
        }
    }
    for (int i = 0; i < n; i++)
        if (s[i] == 'Y') s[i] = 'y';
    return n;
}

}  // namespace en
}  // namespace dcd
