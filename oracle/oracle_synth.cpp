// ORACLE / TEST INFRASTRUCTURE ONLY -- CPU twin of the synthetic corpus generator of BASELINE.json configs 3-5
// (SplitMix64-seeded xoshiro256**, Zipf(s) word ranks, one space between words, ". " after a
// Geometric(mean 12) number of words with the next word capitalised, a 1-6 digit number with p = 0.01, page = words
// until >= min_page_chars characters).  It is a workload definition, not a Docodo function: there is no reference
// file:line to follow.  It exists so that the CPU baseline (bench.py cpu_baseline / --impl reference) indexes exactly
// the bytes the device generator (docodo_b200/csrc/synth.cu) produces; tests/ compares the two byte for byte.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace {
struct Rng {
    uint64_t s[4];
    static uint64_t splitmix(uint64_t& x) {
        uint64_t z = (x += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    void seed(uint64_t x) { for (int i = 0; i < 4; i++) s[i] = splitmix(x); }
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t next() {
        const uint64_t r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

struct Gen {
    const uint8_t* words; const uint32_t* offs; int nwords; std::vector<double> cdf;
    uint64_t seed; int ppd; int minc;
    // appends one document's pages; returns nothing, page lengths pushed to lens
    void doc(uint64_t global_doc, std::vector<uint8_t>& out, std::vector<uint64_t>& lens) const {
        Rng rng;
        rng.seed(seed ^ (global_doc * 0xD6E8FEB86659FD93ull));
        bool cap = true;
        for (int p = 0; p < ppd; p++) {
            const size_t start = out.size();
            while (out.size() - start < (size_t)minc) {
                const double u = rng.uniform();
                if (u < 0.01) {
                    const uint32_t nd = 1 + (uint32_t)(rng.next() % 6);
                    uint64_t v = rng.next();
                    for (uint32_t k = 0; k < nd; k++) { out.push_back((uint8_t)('0' + v % 10)); v /= 10; }
                } else {
                    const double x = rng.uniform();
                    uint32_t lo = 0, hi = (uint32_t)nwords - 1;
                    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (cdf[mid] >= x) hi = mid; else lo = mid + 1; }
                    const uint32_t a = offs[lo], e = offs[lo + 1];
                    for (uint32_t k = a; k < e; k++) {
                        uint8_t ch = words[k];
                        if (cap && k == a && ch >= 'a' && ch <= 'z') ch -= 32;
                        out.push_back(ch);
                    }
                }
                cap = false;
                if (rng.uniform() < 1.0 / 12.0) { out.push_back('.'); cap = true; }
                out.push_back(' ');
            }
            lens.push_back(out.size() - start);
        }
    }
};
struct Corpus { std::vector<uint8_t> text; std::vector<uint64_t> page_offs, doc_first_page; };
}  // namespace

extern "C" {
void* orc_synth_generate(const uint8_t* words, const uint32_t* word_offs, int nwords, double zipf_s, uint64_t seed,
                         int64_t ndocs, int pages_per_doc, int min_page_chars, int64_t first_doc, int nthreads) {
    Gen g{words, word_offs, nwords, {}, seed, pages_per_doc, min_page_chars};
    g.cdf.resize(nwords);
    double h = 0;
    for (int i = 0; i < nwords; i++) h += std::pow((double)(i + 1), -zipf_s);
    double acc = 0;
    for (int i = 0; i < nwords; i++) { acc += std::pow((double)(i + 1), -zipf_s) / h; g.cdf[i] = acc; }
    g.cdf[nwords - 1] = 1.0;
    if (nthreads < 1) nthreads = 1;
    std::vector<std::vector<uint8_t>> parts(nthreads);
    std::vector<std::vector<uint64_t>> lens(nthreads);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++)
        th.emplace_back([&, t] {
            for (int64_t d = ndocs * t / nthreads; d < ndocs * (t + 1) / nthreads; d++) g.doc((uint64_t)(first_doc + d), parts[t], lens[t]);
        });
    for (auto& x : th) x.join();
    Corpus* c = new Corpus();
    c->page_offs.push_back(0);
    for (int t = 0; t < nthreads; t++) {
        c->text.insert(c->text.end(), parts[t].begin(), parts[t].end());
        for (uint64_t l : lens[t]) c->page_offs.push_back(c->page_offs.back() + l);
        std::vector<uint8_t>().swap(parts[t]);
    }
    for (int64_t d = 0; d <= ndocs; d++) c->doc_first_page.push_back((uint64_t)d * pages_per_doc);
    return c;
}
uint64_t orc_synth_text(void* cv, const uint8_t** text) { Corpus* c = (Corpus*)cv; *text = c->text.data(); return c->text.size(); }
uint64_t orc_synth_pages(void* cv, const uint64_t** page_offs, const uint64_t** doc_first_page) {
    Corpus* c = (Corpus*)cv;
    *page_offs = c->page_offs.data();
    *doc_first_page = c->doc_first_page.data();
    return c->page_offs.size() - 1;
}
void orc_synth_free(void* cv) { delete (Corpus*)cv; }
}
