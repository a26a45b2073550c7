// Russian Snowball stemmer, in place on a UTF-16 buffer, host+device.
//
// Replaces Iveonik.Stemmers.RussianStemmer (StemmersNet.Standard 1.1.1; call sites Dict.cs:65, Build.cs:222).
// Written from the published Snowball definition (~2010: vowels "аеиоуыэюя", no yo-normalisation).
// PARITY WITH THE REFERENCE DLL IS UNPINNED (no ru.voc, no test vector in the reference); parity with the
// repo's CPU oracle is tested.  Suffix tables are kept as ASCII literals (unit = 0x0430 + (char - 'A')) so the
// same source compiles for host and device.
#pragma once
#include "common.cuh"
#include "stem_ru_tables.inc"

namespace dcd {
namespace ru {

DCD_HD bool vow(u16 c) {
    return c == 0x0430 || c == 0x0435 || c == 0x0438 || c == 0x043E || c == 0x0443 || c == 0x044B || c == 0x044D ||
           c == 0x044E || c == 0x044F;   // а е и о у ы э ю я
}

// Longest suffix of s[lim..end) found in the '|'-separated list (which is ordered longest first); 0 if none.
DCD_HD int among(const u16* s, int lim, int end, const char* list) {
    int i = 0;
    while (list[i]) {
        int l = 0;
        while (list[i + l] != '|') l++;
        if (end - l >= lim) {
            bool ok = true;
            for (int k = 0; k < l; k++)
                if (s[end - l + k] != (u16)(0x0430 + (list[i + k] - 'A'))) { ok = false; break; }
            if (ok) return l;
        }
        i += l + 1;
    }
    return 0;
}

// Union of a "group 1" list (needs а/я in front, which stays) and a "group 2" list: Snowball's `among` takes the
// longest match over both; returns the new end or -1.
DCD_HD int two_groups(const u16* s, int pV, int end, const char* g1, const char* g2) {
    const int l1 = among(s, pV, end, g1), l2 = among(s, pV, end, g2);
    if (!l1 && !l2) return -1;
    if (l2 >= l1) return end - l2;
    const int st = end - l1;
    if (st - 1 >= pV && (s[st - 1] == 0x0430 || s[st - 1] == 0x044F)) return st;
    return -1;
}

DCD_HD int stem(u16* s, int n) {
    int pV = n, p2 = n;
    {
        int i = 0;
        while (i < n && !vow(s[i])) i++;
        if (i < n) {
            pV = i + 1;
            while (i < n && vow(s[i])) i++;
            if (i < n) {
                i++;
                while (i < n && !vow(s[i])) i++;
                if (i < n) {
                    i++;
                    while (i < n && vow(s[i])) i++;
                    if (i < n) p2 = i + 1;
                }
            }
        }
    }
    int end = n;
    int e = two_groups(s, pV, end, DCD_RU_PG1, DCD_RU_PG2);
    if (e >= 0) {
        end = e;
    } else {
        end -= among(s, pV, end, DCD_RU_REFL);
        int l = among(s, pV, end, DCD_RU_ADJ);
        if (l) {
            end -= l;
            int pe = two_groups(s, pV, end, DCD_RU_PART1, DCD_RU_PART2);
            if (pe >= 0) end = pe;
        } else {
            int ve = two_groups(s, pV, end, DCD_RU_VERB1, DCD_RU_VERB2);
            if (ve >= 0) end = ve;
            else end -= among(s, pV, end, DCD_RU_NOUN);
        }
    }
    if (end > pV && s[end - 1] == 0x0438) end--;                     // и
    {
        int l = among(s, pV, end, DCD_RU_DERIV);
        if (l && end - l >= p2) end -= l;
    }
    {
        auto nn = [&](int e2) { return e2 - 2 >= pV && s[e2 - 1] == 0x043D && s[e2 - 2] == 0x043D; };   // нн
        int l;
        if (nn(end)) end--;
        else if ((l = among(s, pV, end, DCD_RU_SUPER)) != 0) {
            end -= l;
            if (nn(end)) end--;
        } else if (end > pV && s[end - 1] == 0x044C) end--;            // ь
    }
    return end;
}

}  // namespace ru
}  // namespace dcd
